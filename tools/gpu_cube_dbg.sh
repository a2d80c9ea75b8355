#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out
for V in "FLIPB200_FUSED_HALO=0" "FLIPB200_MIXED=0" "FLIPB200_WARM=0" "FLIPB200_FUSED_GATHER=0"; do
  echo "== $V"
  env $V FLIPB200_WARM_DEBUG=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --workload cube-drop --grid-n 512 --steps 2 --warmup 4 > gpurun_out/dbg.json 2> gpurun_out/dbg.err; echo "rc=$?"
  grep "flipb200 solve\|FlipError" gpurun_out/dbg.err | head -6
  python -c "
import json
try:
    d=json.loads(open('gpurun_out/dbg.json').read().strip().splitlines()[-1]); print('ok', d['ms_per_step'], d['checks_ok'])
except Exception as e: print('no line')
"
done
