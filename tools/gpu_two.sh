#!/bin/bash
# 2-GPU check: NCCL slab tests, bench, stage profile.  usage: tools/gpu_two.sh <tag> [gpus]
TAG=${1:-g2}; G=${2:-2}
timeout 600 python -m pytest tests/test_gpu_nccl.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench_$G.json 2> gpurun_out/${TAG}_bench_$G.err; echo "bench$G rc=$?"
grep -v "^\*\|OMP_NUM" gpurun_out/${TAG}_bench_$G.err | tail -3
python - gpurun_out/${TAG}_bench_$G.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k:d.get(k) for k in ['value','ms_per_step','n_gpus']}, 'iteration us', d['kernels']['pcg_iteration']['us'], [s['pcg_iters_total'] for s in d['per_step']], d['per_step'][-1])
PY
FLIPB200_PROFILE_ITER=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_prof_$G.json 2> gpurun_out/${TAG}_prof_$G.err
grep "r0 \|rank 0" gpurun_out/${TAG}_prof_$G.err | head -45
