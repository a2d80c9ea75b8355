#!/bin/bash
TAG=${1:-hd}; G=${2:-2}
FLIPB200_HALO_DEBUG=1 FLIPB200_PROFILE_ITER=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 1 --warmup 1 --no-cpu --no-e2e --kernel-iters 20 > gpurun_out/${TAG}_prof_$G.json 2> gpurun_out/${TAG}_prof_$G.err
grep "r0 \|rank 0\|h0 " gpurun_out/${TAG}_prof_$G.err | head -90
