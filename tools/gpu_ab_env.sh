#!/bin/bash
# A/B of environment switches on one GPU: usage  gpurun -- 'bash tools/gpu_ab_env.sh <tag> "A=1" "FLIPB200_CHEB=20,0.005" ...'
TAG=$1; shift
mkdir -p gpurun_out
n=0
for V in "$@"; do
  n=$((n+1)); OUT=gpurun_out/${TAG}_v$n
  env $V timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra > $OUT.json 2> $OUT.err
  python - <<PY
import json
d=json.loads(open('$OUT.json').read().strip().splitlines()[-1])
b=d['roofline']['by_kernel']
print('%-28s ms/step %.2f iter %.1f tail %.1f small-levels %.1f iters %s ok %s' % ('$V', d['ms_per_step'], b['pcg_iteration']['us'], b.get('pcg/tail (small levels + coarsest)',{}).get('us_per_iteration',-1), b['pcg_iteration']['us_levels_below_0'], [s['pcg_iters_total'] for s in d['per_step']], d['checks_ok']))
PY
done
