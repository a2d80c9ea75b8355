#!/bin/bash
TAG=$1; G=${2:-4}
run() { local name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 6 --warmup 3 --no-cpu --no-e2e --kernel-iters 10 > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$?"
  python - gpurun_out/${TAG}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d['ms_per_step'], d['kernels']['pcg_iteration']['us'], [(round(s['ms_total'],1), round(s['ms_pcg'],1), s['pcg_iters_total']) for s in d['per_step']], d['clocks'])
except Exception as e: print('ERR',e)
PY
}
run smi0 A=1
run nosmi FLIPB200_BENCH_NOSMI=1
run nosmi_nograph FLIPB200_BENCH_NOSMI=1 FLIPB200_NO_GRAPH=1
