#!/bin/bash
# sweep of the restrict / prolong z-chunk on one GPU (stage profile of the PCG iteration)
TAG=$1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
for V in "A=1" "FLIPB200_RS_ZC=4 FLIPB200_PR_ZC=4" "FLIPB200_RS_ZC=16 FLIPB200_PR_ZC=16" "FLIPB200_RS_ZC=2 FLIPB200_PR_ZC=2"; do
  echo "== $V"
  env $V FLIPB200_PROFILE_ITER=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_p1.json 2> gpurun_out/${TAG}_p1.err
  grep "total\|restrict\|prolong" gpurun_out/${TAG}_p1.err
  python -c "
import json
d=json.loads(open('gpurun_out/${TAG}_p1.json').read().strip().splitlines()[-1]); print('graph iteration us', d['kernels']['pcg_iteration']['us'], [s['pcg_iters_total'] for s in d['per_step']])"
done
