#!/bin/bash
TAG=$1; G=${2:-2}
timeout 300 python -m pytest tests/test_gpu_nccl.py -x -q 2>&1 | tail -3
run() { local name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 3 --warmup 3 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$?"; grep -v "^\*\|OMP_NUM" gpurun_out/${TAG}_$name.err | tail -3
  python - gpurun_out/${TAG}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d['ms_per_step'], d['value'], d['kernels']['pcg_iteration']['us'], d['config'].get('engine'), [(round(s['ms_total'],1), s['pcg_iters_total']) for s in d['per_step']])
except Exception as e: print('ERR',e)
PY
}
run fusedred A=1
run ncclred FLIPB200_FUSED_REDUCE=0
