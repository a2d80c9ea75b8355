#!/bin/bash
# multi-GPU round: NCCL slab test + scaling bench.  usage: tools/gpu_multi.sh <tag> <ngpu> [n]
TAG=${1:-m}; NG=${2:-2}; N=${3:-256}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
if [ "${SKIP_TEST:-0}" = "0" ]; then timeout 600 python -m pytest tests/test_gpu_nccl.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${TAG}_pytest.log; fi
timeout 300 python bench.py --grid-n $N --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench_1.json 2> gpurun_out/${TAG}_bench_1.err; echo "bench1 rc=$?"
G=2
while [ $G -le $NG ]; do
  NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --grid-n $N --gpus $G --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench_$G.json 2> gpurun_out/${TAG}_bench_$G.err; echo "bench$G rc=$?"
  tail -3 gpurun_out/${TAG}_bench_$G.err
  G=$((G*2))
done
for f in gpurun_out/${TAG}_bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ['value','ms_per_step','n_gpus']}, d['e2e'] and d['e2e']['value'], d['kernels'], d['per_step'][-1])
except Exception as e: print('ERR',e)
PY
done
