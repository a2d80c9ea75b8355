#!/bin/bash
# One gpurun call: GPU parity tests, bench, ncu launch list, ncu --set full of the hot kernels.
# usage: tools/gpu_round.sh <tag>
set -u
TAG=${1:-r1}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err; echo "ref rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --kernel-iters 3"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "launch list rc=$?"
# level-0 launches: the first launch of the V-cycle kernels is level 0, the 4th prolong of a cycle is level 1 -> 0
for KS in k_spmv:2 k_update_xr:2 k_update_p:2 k_mg_smooth:0 k_mg_resid:0 k_mg_restrict:0 k_mg_prolong:3 k_mg_tail:0 k_p2g_cell:1 k_g2p_cell:0 k_advect:0 k_reorder:1 k_key_count:1 k_classify:1 k_rhs_div:1 k_vel_update:1 k_cell_delta:0; do
  K=${KS%%:*}; SK=${KS##*:}
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^${K}" -s $SK -c 1 -f -o gpurun_out/${TAG}_${K} $CMD > gpurun_out/${TAG}_ncu_${K}.log 2>&1
  echo "ncu $K rc=$?"
done
ls -la gpurun_out | head -60
