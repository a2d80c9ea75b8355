#!/bin/bash
TAG=${1:-r2f}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest default rc=$?"; tail -8 gpurun_out/${TAG}_pytest.log
FLIPB200_WARM_DEBUG=1 timeout 300 python bench.py --grid-n 128 --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra 2>&1 >/dev/null | grep "flipb200 solve" | tail -16
FLIPB200_WARM_DEBUG=1 timeout 300 python bench.py --workload dam-break-column --grid-n 128 --steps 6 --warmup 10 --no-cpu --no-e2e --no-extra 2>&1 >/dev/null | grep "flipb200 solve" | tail -12
for V in "A=0"; do
  echo "== $V"
  env $V timeout 400 python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_bench.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
    print('  value %.4g  ms/step %.2f checks %s' % (d['value'], d['ms_per_step'], d['checks_ok']))
    for s in d['per_step']: print('  ', s)
except Exception as e:
    print('  parse failed', e)
PY
done
