#!/bin/bash
# warm start + mixed-precision defaults: full GPU suite, A/B of the switches
TAG=${1:-r2d}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest default rc=$?"; tail -15 gpurun_out/${TAG}_pytest.log
n=0
for V in "A=0" "FLIPB200_WARM=0" "FLIPB200_MIXED=0" "FLIPB200_WARM=0 FLIPB200_MIXED=0"; do
  n=$((n+1))
  echo "== variant $n: $V"
  env $V timeout 400 python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra > gpurun_out/${TAG}_v${n}_bench.json 2> gpurun_out/${TAG}_v${n}_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_v${n}_bench.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_v${n}_bench.json').read().strip().splitlines()[-1])
    print('  value %.4g  ms/step %.2f checks %s iter %.1f us' % (d['value'], d['ms_per_step'], d['checks_ok'], d['roofline']['by_kernel']['pcg_iteration']['us']))
    for s in d['per_step']: print('  ', s)
except Exception as e:
    print('  parse failed', e)
PY
done
for W in dam-break-column cube-drop; do
for V in "A=0" "FLIPB200_WARM=0"; do
  echo "== $W $V"
  env $V timeout 400 python bench.py --workload $W --grid-n $([ $W = cube-drop ] && echo 512 || echo 256) --steps 8 --warmup 4 --no-cpu --no-e2e > gpurun_out/${TAG}_w_bench.json 2> gpurun_out/${TAG}_w_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_w_bench.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_w_bench.json').read().strip().splitlines()[-1])
    print('  value %.4g  ms/step %.2f checks %s' % (d['value'], d['ms_per_step'], d['checks_ok']))
    for s in d['per_step']: print('  ', s)
except Exception as e:
    print('  parse failed', e)
PY
done
done
