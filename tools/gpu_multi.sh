#!/bin/bash
# usage: gpurun --gpus N --timeout 1500 -- 'bash tools/gpu_multi.sh <tag> <N> [steps] [extra bench args]'
TAG=${1:-r2m}; N=${2:-2}; STEPS=${3:-5}; shift 3
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps $STEPS --warmup 3 "$@" > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err; echo "bench N=$N rc=$?"
tail -c 1500 gpurun_out/${TAG}_bench_n${N}.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_bench_n${N}.json').read().strip().splitlines()[-1])
    print('value %.4g ms/step %.2f e2e %s checks_ok %s' % (d['value'], d['ms_per_step'], d.get('e2e') and '%.4g' % d['e2e']['value'], d['checks_ok']))
    print('check', json.dumps(d['check'])[:1200])
    print('iter us', d['roofline']['by_kernel']['pcg_iteration']['us'])
    for k in ('pcg','dam_break_column','cube_drop_512'):
        if k in d: print(k, 'value %.4g ms/step %.2f' % (d[k]['value'], d[k]['ms_per_step']), json.dumps(d[k]['check'])[:1400])
    print([ (s['passes'], s['pcg_iters_total'], round(s['ms_total'],1)) for s in d['per_step']])
except Exception as e:
    print('parse failed', e)
PY
