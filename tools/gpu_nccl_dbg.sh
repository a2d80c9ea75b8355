#!/bin/bash
# 2 GPUs: the NCCL slab test with halo copies over NCCL (param [2]) under switch combinations
mkdir -p gpurun_out
for V in "A=1" "FLIPB200_WARM=0" "FLIPB200_NO_GRAPH=1" "FLIPB200_NO_GRAPH=1 FLIPB200_WARM=0" "FLIPB200_FUSED_REDUCE=0 FLIPB200_FUSED_GATHER=0"; do
  env $V timeout 300 python -m pytest "tests/test_gpu_nccl.py::test_nccl_slabs_match_single[2]" -x -q > "gpurun_out/nccl_dbg_$V.log" 2>&1; echo "$V rc=$?"
  grep "AssertionError\|passed\|failed" "gpurun_out/nccl_dbg_$V.log" | head -4
done
