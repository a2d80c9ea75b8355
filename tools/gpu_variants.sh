#!/bin/bash
# 2-GPU variants of the step: graph x fused-halo.  usage: tools/gpu_variants.sh <tag> <ngpu>
TAG=$1; G=${2:-2}
run() { # name env...
  local name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 2 --warmup 2 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$?"
  python - gpurun_out/${TAG}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d['ms_per_step'], d['kernels']['pcg_iteration']['us'], d['config'].get('engine'), d['per_step'][-1]['pcg_iters_total'])
except Exception as e: print('ERR',e)
PY
}
run graph_fused A=1
run nograph_fused FLIPB200_NO_GRAPH=1
run graph_nccl FLIPB200_FUSED_HALO=0
run nograph_nccl FLIPB200_NO_GRAPH=1 FLIPB200_FUSED_HALO=0
