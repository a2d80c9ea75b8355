#!/bin/bash
# N-GPU check (default 4): NCCL slab tests (world = min(N, 4)) and one bench line
TAG=${1:-g4}; G=${2:-4}
timeout 600 python -m pytest tests/test_gpu_nccl.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench_$G.json 2> gpurun_out/${TAG}_bench_$G.err; echo "bench$G rc=$?"
grep -v "^\*\|OMP_NUM" gpurun_out/${TAG}_bench_$G.err | tail -3
python - gpurun_out/${TAG}_bench_$G.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k:d.get(k) for k in ['value','ms_per_step','n_gpus']}, 'iteration us', d['kernels']['pcg_iteration']['us'], d['config'].get('engine'), [s['pcg_iters_total'] for s in d['per_step']])
PY
