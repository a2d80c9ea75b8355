#!/bin/bash
# 1-GPU check of a kernel change: GPU tests, stage profile of one iteration, short bench.
TAG=$1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
FLIPB200_PROFILE_ITER=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_p1.json 2> gpurun_out/${TAG}_p1.err; echo rc=$?
grep -A40 "PCG iteration profile" gpurun_out/${TAG}_p1.err | head -45
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['kernels']['pcg_iteration'], [s['pcg_iters_total'] for s in d['per_step']])
PY
