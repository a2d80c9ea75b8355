#!/bin/bash
# A/B of env-switchable variants on one GPU: tools/gpu_ab.sh TAG "A=1" "FLIPB200_X=0" ...
# Runs the GPU tests once (default variant), then per variant the stage profile of one PCG
# iteration and a 3-step bench.
TAG=$1; shift
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
n=0
for V in "$@"; do
  n=$((n+1))
  echo "== variant $n: $V"
  env $V FLIPB200_PROFILE_ITER=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_v${n}_p.json 2> gpurun_out/${TAG}_v${n}_p.err
  grep -A40 "PCG iteration profile" gpurun_out/${TAG}_v${n}_p.err | grep "${AB_GREP:-total}"
  env $V timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_v${n}_bench.json 2> gpurun_out/${TAG}_v${n}_bench.err; echo "bench rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_v${n}_bench.json').read().strip().splitlines()[-1])
print('  value %.4g  ms/step %.2f  iteration %.1f us  iters/step %s  p2g %.2f ms' % (d['value'], d['ms_per_step'], d['kernels']['pcg_iteration']['us'], [s['pcg_iters_total'] for s in d['per_step']], d['kernels']['p2g']['ms']))
print('  per_step', {k: round(v, 2) for k, v in d['per_step'][-1].items() if k.startswith('ms_')})
PY
done
