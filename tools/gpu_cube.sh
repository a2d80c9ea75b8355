#!/bin/bash
# usage: gpurun --gpus N -- 'bash tools/gpu_cube.sh <tag> <N> [extra args]'
TAG=${1:-cube}; N=${2:-2}; shift 2
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload cube-drop --grid-n 512 --steps 10 --warmup 4 "$@" > gpurun_out/${TAG}_cube_n${N}.json 2> gpurun_out/${TAG}_cube_n${N}.err; echo "cube N=$N rc=$?"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/${TAG}_cube_n${N}.err | tail -5
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_cube_n${N}.json').read().strip().splitlines()[-1])
    print('value %.4g ms/step %.2f checks %s' % (d['value'], d['ms_per_step'], d['checks_ok']), d['config'].get('slab_boundaries'))
    print(json.dumps(d['check'])[:900])
    for s in d['per_step'][-3:]: print('  ', s)
except Exception as e:
    print('parse failed', e)
PY
