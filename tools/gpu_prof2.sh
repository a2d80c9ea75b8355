#!/bin/bash
# usage: gpurun --gpus G -- 'bash tools/gpu_prof2.sh <tag> <G>'   stage profile of one PCG iteration on G GPUs, with variants
TAG=$1; G=${2:-2}
mkdir -p gpurun_out
for V in "A=1" "FLIPB200_HALO_ACQ=gpu" "FLIPB200_HALO_DEBUG=1"; do
env $V FLIPB200_PROFILE_ITER=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 3 --warmup 3 --no-cpu --no-e2e --no-extra --kernel-iters 20 > gpurun_out/${TAG}_p${G}_$V.json 2> gpurun_out/${TAG}_p${G}_$V.err; echo "$V rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_p${G}_$V.json').read().strip().splitlines()[-1])
print('$V', 'ms/step %.2f iter us %.1f' % (d['ms_per_step'], d['roofline']['by_kernel']['pcg_iteration']['us']))
PY
done
grep "h0 \|halo timestamps" gpurun_out/${TAG}_p${G}_FLIPB200_HALO_DEBUG=1.err | head -60
