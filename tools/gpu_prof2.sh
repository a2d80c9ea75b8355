#!/bin/bash
TAG=$1; G=${2:-2}
FLIPB200_PROFILE_ITER=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_p1.json 2> gpurun_out/${TAG}_p1.err; echo rc=$?
grep -A40 "PCG iteration profile" gpurun_out/${TAG}_p1.err | head -45
for V in "A=1" "FLIPB200_FUSED_HALO=0"; do
env $V FLIPB200_PROFILE_ITER=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_p$G.json 2> gpurun_out/${TAG}_p${G}_$V.err; echo rc=$?
grep "r0 \|rank 0" gpurun_out/${TAG}_p${G}_$V.err | head -45
done
