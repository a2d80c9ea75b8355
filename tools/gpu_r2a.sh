#!/bin/bash
# Round-2 opener (one GPU): A/B of the variants written at the end of round 1 + the kbench notebook.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_r2a.sh r2a'
TAG=${1:-r2a}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 300 ./tools/kbench > gpurun_out/${TAG}_kbench.txt 2>&1; echo "kbench rc=$?"
FLIPB200_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "p2g_classify or g2p_advect" > gpurun_out/${TAG}_pytest_exp.log 2>&1; echo "pytest experimental params rc=$?"; tail -5 gpurun_out/${TAG}_pytest_exp.log
n=0
for V in "A=0" "FLIPB200_P2G=v4" "FLIPB200_P2G=v4c" "FLIPB200_G2P=v4" "FLIPB200_G2P=v4c" "FLIPB200_G2P=f1" "FLIPB200_G2P=f4" "FLIPB200_SORT=k" "FLIPB200_P2G=v4 FLIPB200_G2P=f4 FLIPB200_SORT=k" "FLIPB200_UP4=1" "FLIPB200_MG_H16=1" "FLIPB200_P32=1" "FLIPB200_MG_H16=1 FLIPB200_P32=1 FLIPB200_UP4=1"; do
  n=$((n+1))
  echo "== variant $n: $V"
  env $V timeout 400 python bench.py --steps 4 --warmup 4 --no-cpu --no-e2e > gpurun_out/${TAG}_v${n}_bench.json 2> gpurun_out/${TAG}_v${n}_bench.err; echo "bench rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_v${n}_bench.json').read().strip().splitlines()[-1])
    k=d['kernels']
    print('  value %.4g  ms/step %.2f  p2g %.2f ms (%.2f)  g2p+advect %.2f ms (%.2f)  particle phase %.2f ms (%.2f)  iter %.1f us  spmv %.1f us' % (
        d['value'], d['ms_per_step'], k['p2g']['ms'], k['p2g']['frac'], k['g2p_advect']['ms'], k['g2p_advect']['frac'],
        k['particle_phase']['ms'], k['particle_phase']['frac'], k['pcg_iteration']['us'], d['roofline']['us_per_launch']))
    for s in d['per_step']: print('  ', {a: round(v, 2) for a, v in s.items()})
except Exception as e:
    print('  parse failed', e)
PY
done
