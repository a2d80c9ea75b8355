#!/bin/bash
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_r2b.sh r2b'
TAG=${1:-r2b}
mkdir -p gpurun_out
timeout 300 ./tools/kbench 256 1 > gpurun_out/${TAG}_kbench_stencil.txt 2>&1; echo "kbench rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
print('value %.4g ms/step %.2f e2e %.4g checks_ok %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['checks_ok']))
print('check', json.dumps(d['check'])[:900])
print('dominant', d['roofline']['dominant'])
for k,v in d['roofline']['by_kernel'].items(): print('  ', k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a!='what'})
for k in ('pcg','dam_break_column','cube_drop_512','grid_128','default_scene'):
    if k in d: print(k, json.dumps({a:b for a,b in d[k].items() if a not in ('per_step','config')})[:1200])
print('cpu', json.dumps(d['cpu_baseline'])[:1500])
PY
