#!/bin/bash
# One gpurun call (one GPU): launch list + two `ncu --set full` captures of the current defaults on the 256^3 headline scene.
#   A: particle / grid kernels of the first step     B: three PCG iterations, every level (tools/ncu_traffic.py keeps level 0)
# usage: tools/gpu_ncu.sh <tag>
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-extra --kernel-iters 3"
timeout 300 $CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none \
  -k regex:"^(k_key_count|k_scan|k_reorder|k_p2g|k_classify|k_rhs|k_g2p|k_speed|k_advect|k_vel_update|k_cell_delta|k_relax|k_warm)" \
  -c 32 -f -o gpurun_out/${TAG}_k_particles $CMD > gpurun_out/${TAG}_ncu_A.log 2>&1
echo "ncu A rc=$?"
timeout 900 ncu --set full --clock-control none \
  -k regex:"^(k_spmv|k_update|k_mg_|k_pcg)" -s 90 -c 40 -f -o gpurun_out/${TAG}_k_pcg $CMD > gpurun_out/${TAG}_ncu_B.log 2>&1
echo "ncu B rc=$?"
# the reports are too large to travel back (64 MiB limit): reduce them here and keep only the tables
for R in particles pcg; do ncu -i gpurun_out/${TAG}_k_$R.ncu-rep --page raw --csv > gpurun_out/${TAG}_k_${R}_raw.csv 2>/dev/null; done
python tools/ncu_traffic.py ${TAG} gpurun_out/${TAG}_ncu_summary.txt --box
rm -f gpurun_out/${TAG}_k_*.ncu-rep
du -sh gpurun_out; ls -la gpurun_out | head -30
