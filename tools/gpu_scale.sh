#!/bin/bash
# weak-scaling round on one box: tools/gpu_scale.sh <tag> <max gpus>
TAG=${1:-sc}; NG=${2:-8}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_nccl.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench_1.json 2> gpurun_out/${TAG}_bench_1.err; echo "bench1 rc=$?"
G=2
while [ $G -le $NG ]; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench_$G.json 2> gpurun_out/${TAG}_bench_$G.err; echo "bench$G rc=$?"
  grep -v "^\*\|OMP_NUM" gpurun_out/${TAG}_bench_$G.err | tail -3
  G=$((G*2))
done
FLIPB200_PROFILE_ITER=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $NG --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_prof_$NG.json 2> gpurun_out/${TAG}_prof_$NG.err
grep "r0 \|rank 0" gpurun_out/${TAG}_prof_$NG.err | head -45
for f in gpurun_out/${TAG}_bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ['value','ms_per_step','n_gpus']}, 'iteration us', d['kernels']['pcg_iteration']['us'], d['config'].get('engine'), [s['pcg_iters_total'] for s in d['per_step']], d['per_step'][-1])
except Exception as e: print('ERR',e)
PY
done
