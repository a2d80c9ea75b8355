#!/bin/bash
# Round-2 opener (one gpurun call, one GPU): the particle-kernel variants written after the GPU budget
# of round 1 was spent.  1. parity tests of the default build; 2. the same tests with every
# experimental variant enabled as an extra test parameter; 3. the whole GPU suite with the variants
# forced on through the environment (multi-step scenes, bet won and lost); 4. 3-step bench per variant.
# usage: gpurun --timeout 2400 -- 'bash tools/gpu_r2_particles.sh r2a'
TAG=${1:-r2a}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest default rc=$?"; tail -2 gpurun_out/${TAG}_pytest.log
FLIPB200_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "p2g_classify or g2p_advect" > gpurun_out/${TAG}_pytest_exp.log 2>&1; echo "pytest experimental params rc=$?"; tail -5 gpurun_out/${TAG}_pytest_exp.log
for V in "FLIPB200_P2G=v4 FLIPB200_G2P=v4" "FLIPB200_P2G=v4 FLIPB200_G2P=f4" "FLIPB200_G2P=f1" "FLIPB200_SORT=k" "FLIPB200_SORT=k FLIPB200_G2P=f4"; do
  T=$(echo "$V" | tr ' =' '__')
  env $V timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -q > gpurun_out/${TAG}_pytest_${T}.log 2>&1; echo "pytest [$V] rc=$?"; tail -3 gpurun_out/${TAG}_pytest_${T}.log
done
n=0
for V in "A=0" "FLIPB200_P2G=v4" "FLIPB200_P2G=v4c" "FLIPB200_G2P=v4" "FLIPB200_G2P=v4c" "FLIPB200_G2P=f1" "FLIPB200_G2P=f4" "FLIPB200_G2P=f4c" "FLIPB200_SORT=k" "FLIPB200_P2G=v4 FLIPB200_G2P=f4" "FLIPB200_P2G=v4 FLIPB200_G2P=f4 FLIPB200_SORT=k"; do
  n=$((n+1))
  echo "== variant $n: $V"
  env $V timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_v${n}_bench.json 2> gpurun_out/${TAG}_v${n}_bench.err; echo "bench rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_v${n}_bench.json').read().strip().splitlines()[-1])
k=d['kernels']
print('  value %.4g  ms/step %.2f  p2g %.2f ms (%.2f)  g2p+advect %.2f ms (%.2f)  particle phase %.2f ms (%.2f)' % (
    d['value'], d['ms_per_step'], k['p2g']['ms'], k['p2g']['frac'], k['g2p_advect']['ms'], k['g2p_advect']['frac'],
    k['particle_phase']['ms'], k['particle_phase']['frac']))
print('  per_step', {a: round(v, 2) for a, v in d['per_step'][-1].items() if a.startswith('ms_')})
PY
done
# PCG-side experiments: four-node direction update, fp16 storage of b0 / the level-0 residual, fp32 search direction
for V in "FLIPB200_UP4=1" "FLIPB200_MG_H16=1" "FLIPB200_P32=1" "FLIPB200_MG_H16=1 FLIPB200_P32=1"; do
  T=$(echo "$V" | tr ' =' '__')
  env $V timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -q > gpurun_out/${TAG}_pytest_${T}.log 2>&1; echo "pytest [$V] rc=$?"; tail -3 gpurun_out/${TAG}_pytest_${T}.log
  env $V timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_${T}_bench.json 2> gpurun_out/${TAG}_${T}_bench.err; echo "bench [$V] rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_${T}_bench.json').read().strip().splitlines()[-1])
print('  value %.4g  ms/step %.2f  iteration %.1f us  iters/step %s' % (d['value'], d['ms_per_step'], d['kernels']['pcg_iteration']['us'], [s['pcg_iters_total'] for s in d['per_step']]))
PY
done
