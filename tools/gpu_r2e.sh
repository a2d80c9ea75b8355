#!/bin/bash
TAG=${1:-r2e}
mkdir -p gpurun_out
FLIPB200_WARM_DEBUG=1 timeout 300 python bench.py --grid-n 128 --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra 2>&1 >/dev/null | grep "flipb200 solve" | head -40
FLIPB200_WARM_DEBUG=1 timeout 300 python bench.py --workload cube-drop --grid-n 256 --steps 4 --warmup 4 --no-cpu --no-e2e --no-extra 2>&1 >/dev/null | grep "flipb200 solve" | head -40
for V in "A=0" "FLIPB200_G2P=fp" "FLIPB200_G2P=fp FLIPB200_SORT=k" "FLIPB200_G2P=p"; do
  echo "== $V"
  env $V timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e --no-extra > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_bench.err
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
FLIPB200_G2P=fp timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -5
FLIPB200_G2P=fp FLIPB200_SORT=k timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -5
