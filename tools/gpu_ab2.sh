#!/bin/bash
# like gpu_ab.sh without the tests and the stage profile: 3-step bench per variant
TAG=$1; shift
n=0
for V in "$@"; do
  n=$((n+1))
  echo "== variant $n: $V"
  env $V timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_v${n}_bench.json 2> gpurun_out/${TAG}_v${n}_bench.err; echo "bench rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_v${n}_bench.json').read().strip().splitlines()[-1])
print('  value %.4g  ms/step %.2f  iteration %.1f us  iters/step %s  p2g %.2f ms' % (d['value'], d['ms_per_step'], d['kernels']['pcg_iteration']['us'], [s['pcg_iters_total'] for s in d['per_step']], d['kernels']['p2g']['ms']))
print('  per_step', {k: round(v, 2) for k, v in d['per_step'][-1].items() if k.startswith('ms_')})
PY
done
