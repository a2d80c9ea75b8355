#!/bin/bash
# one short A/B of two library builds on the headline scene: step time and the per-phase device times
mkdir -p gpurun_out
for SFX in "$@"; do
  [ "$SFX" = "-" ] && LIBP="" || LIBP="$PWD/fluid-simulation_b200/libflipb200$SFX.so"
  FLIPB200_LIB_PATH=$LIBP timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra --kernel-iters 5 > gpurun_out/q$SFX.json 2> gpurun_out/q$SFX.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/q$SFX.json').read().strip().splitlines()[-1])
s=d['per_step']
print('lib$SFX ms/step %.3f ok %s' % (d['ms_per_step'], d['checks_ok']), ' '.join('%s %.3f' % (k, sum(x[k] for x in s[:3])/3) for k in ('ms_sort','ms_p2g','ms_grid','ms_pcg','ms_g2p')))
PY
done
