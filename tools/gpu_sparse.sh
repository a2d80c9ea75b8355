#!/bin/bash
# one GPU: GPU suite with the active-region sweeps (default), then the cube drop in a 512^3 box and the column dam break with and without them
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/sp_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/sp_pytest.log
for V in 1 0; do
  FLIPB200_SPARSE=$V timeout 200 python bench.py --workload cube-drop --grid-n 512 --steps 8 --warmup 4 > gpurun_out/sp_cube_$V.json 2> gpurun_out/sp_cube_$V.err; echo "cube SPARSE=$V rc=$?"
  FLIPB200_SPARSE=$V timeout 200 python bench.py --workload dam-break-column --steps 6 --warmup 3 --no-e2e --no-cpu > gpurun_out/sp_col_$V.json 2> gpurun_out/sp_col_$V.err; echo "column SPARSE=$V rc=$?"
  python - <<PY
import json
for w in ('cube','col'):
    d=json.loads(open('gpurun_out/sp_%s_$V.json' % w).read().strip().splitlines()[-1])
    print('SPARSE=$V', w, 'ms/step %.2f value %.4g checks %s' % (d['ms_per_step'], d['value'], d.get('checks_ok')), [ (round(s['ms_sort'],2), round(s['ms_p2g'],2), round(s['ms_grid'],2), round(s['ms_pcg'],2), round(s['ms_g2p'],2)) for s in d['per_step'][-2:]])
PY
done
timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extra > gpurun_out/sp_head.json 2> gpurun_out/sp_head.err; echo "headline rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/sp_head.json').read().strip().splitlines()[-1])
print('headline ms/step %.2f value %.4g checks %s' % (d['ms_per_step'], d['value'], d['checks_ok']))
PY
