#!/bin/bash
TAG=${1:-r2g}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest default rc=$?"; tail -12 gpurun_out/${TAG}_pytest.log
for V in "A=0" "FLIPB200_P2G=cell FLIPB200_G2P=cell" "FLIPB200_P2G=cell" "FLIPB200_G2P=cell"; do
  echo "== $V"
  env $V timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e --no-extra > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_bench.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
    print('  value %.4g  ms/step %.2f checks %s' % (d['value'], d['ms_per_step'], d['checks_ok']))
    k=d['roofline']['by_kernel']
    print('  p2g %.2f ms (%.2f) g2p %.2f (%.2f) sort %.2f phase %.2f (%.2f)' % (k['p2g']['ms'], k['p2g']['frac'], k['g2p_advect']['ms'], k['g2p_advect']['frac'], k['sort']['ms'], k['particle_phase']['ms'], k['particle_phase']['frac']))
    for s in d['per_step']: print('  ', s)
except Exception as e:
    print('  parse failed', e)
PY
done
for W in dam-break-column cube-drop; do
for V in "A=0" "FLIPB200_P2G=cell FLIPB200_G2P=cell"; do
  echo "== $W $V"
  env $V timeout 400 python bench.py --workload $W --grid-n $([ $W = cube-drop ] && echo 512 || echo 256) --steps 6 --warmup 4 --no-cpu --no-e2e > gpurun_out/${TAG}_w_bench.json 2> gpurun_out/${TAG}_w_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_w_bench.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_w_bench.json').read().strip().splitlines()[-1])
    print('  value %.4g  ms/step %.2f checks %s' % (d['value'], d['ms_per_step'], d['checks_ok']))
    for s in d['per_step'][-3:]: print('  ', s)
except Exception as e:
    print('  parse failed', e)
PY
done
done
