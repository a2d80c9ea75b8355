#!/bin/bash
# A/B of library builds (fluid-simulation_b200/build.py with FLIPB200_BUILD_SUFFIX / FLIPB200_BUILD_DEFS) on G GPUs
# usage: gpurun [--gpus G] -- 'bash tools/gpu_ab_lib.sh <tag> <G> "<suffix list, - for the product>"'
TAG=$1; G=${2:-1}; shift 2
mkdir -p gpurun_out
for round in 1 2; do
for SFX in "$@"; do
  [ "$SFX" = "-" ] && LIBP="" || LIBP="$PWD/fluid-simulation_b200/libflipb200$SFX.so"
  OUT=gpurun_out/${TAG}_g${G}_lib${SFX}_$round
  if [ "$G" = "1" ]; then
    FLIPB200_LIB_PATH=$LIBP timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra > $OUT.json 2> $OUT.err
  else
    FLIPB200_LIB_PATH=$LIBP timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $G --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra > $OUT.json 2> $OUT.err
  fi
  python - <<PY
import json
d=json.loads(open('$OUT.json').read().strip().splitlines()[-1])
b=d['roofline']['by_kernel']
g=lambda k: b.get(k,{}).get('us_per_iteration',float('nan'))
print('lib$SFX G=$G ms/step %.2f iter %.1f | spmv %.1f xr %.1f sm %.1f rs %.1f rst %.1f pro %.1f sm1 %.1f sm2 %.1f up %.1f | small-levels %.1f | ok %s' % (d['ms_per_step'], b['pcg_iteration']['us'], g('pcg/k_spmv'), g('pcg/k_update_xr'), g('pcg/L0 halo+smooth'), g('pcg/L0 halo+resid'), g('pcg/L0 halo+restrict'), g('pcg/L0 halo+prolong'), g('pcg/L0 halo+smooth (post 1)'), g('pcg/L0 halo+smooth (post 2)'), g('pcg/k_update_p'), b['pcg_iteration']['us_levels_below_0'], d['checks_ok']))
PY
done; done
