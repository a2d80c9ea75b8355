#!/bin/bash
# one GPU: the whole GPU suite + the default bench line (what the driver runs at round end)
TAG=${1:-full}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps ${2:-6} --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
print('value %.4g ms/step %.2f e2e %.4g (%.1f ms) checks_ok %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['checks_ok']))
print('dominant', d['roofline']['dominant'])
for k,v in d['roofline']['by_kernel'].items():
    if 'frac' in v or k in ('sort','pcg_iteration'): print('  ', k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a not in ('what','unit')})
for k in ('pcg','dam_break_column','cube_drop_512','grid_128','default_scene','fp32_mode'):
    if k in d: print(k, 'value %.4g ms/step %.2f' % (d[k]['value'], d[k]['ms_per_step']), d[k].get('check',{}).get('ok'))
print('cpu', {k:v for k,v in d['cpu_baseline'].items() if k not in ('sample','same_config_gpu')}, d['cpu_baseline']['same_config_gpu']['value'])
for s in d['per_step']: print('  ', s)
PY
