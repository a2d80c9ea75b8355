#!/bin/bash
TAG=${1:-r2h}
mkdir -p gpurun_out
for V in "A=0" "FLIPB200_P2G=cell FLIPB200_G2P=cell"; do
  echo "== $V"
  env $V timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e --no-extra > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_bench.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
    print('  value %.4g  ms/step %.2f checks %s' % (d['value'], d['ms_per_step'], d['checks_ok']))
    k=d['roofline']['by_kernel']
    print('  p2g %.2f ms (%.2f) g2p %.2f (%.2f) sort %.2f phase %.2f (%.2f)' % (k['p2g']['ms'], k['p2g']['frac'], k['g2p_advect']['ms'], k['g2p_advect']['frac'], k['sort']['ms'], k['particle_phase']['ms'], k['particle_phase']['frac']))
    for s in d['per_step'][-2:]: print('  ', s)
except Exception as e:
    print('  parse failed', e)
PY
done
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-extra --kernel-iters 3"
for K in k_p2g_rows k_g2p_advect_rows; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^${K}" -s 1 -c 1 -f -o gpurun_out/${TAG}_${K} $CMD > gpurun_out/${TAG}_ncu_${K}.log 2>&1
  echo "ncu $K rc=$?"
done
