#!/bin/bash
TAG=${1:-s8}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_nccl.py -x -q -k "single[0]" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
for G in 8 4; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench_$G.json 2> gpurun_out/${TAG}_bench_$G.err; echo "bench$G rc=$?"
  grep -v "^\*\|OMP_NUM" gpurun_out/${TAG}_bench_$G.err | tail -5
  python - gpurun_out/${TAG}_bench_$G.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ['value','ms_per_step','n_gpus']}, d['e2e'] and d['e2e']['value'], d['kernels']['pcg_iteration']['us'], d['config'].get('engine'), d['per_step'][-1])
except Exception as e: print('ERR',e)
PY
done
