#!/bin/bash
# ncu captures of this round's kernels (one GPU; run under gpurun AFTER the plain bench exited 0):
#   bash profiles/ncu_capture.sh
set -x
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_plain.log 2>&1 || exit 1
NCU="ncu --set full --clock-control none --import-source on"
# K4 (one launch of the bench workload), K1, K3
$NCU -k regex:dipole_sym_kernel -s 12 -c 1 -o gpurun_out/r01_k4_v8 -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_k4.log 2>&1
$NCU -k regex:pair_cluster_kernel -s 1 -c 1 -o gpurun_out/r01_k1_cluster -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_k1.log 2>&1
$NCU -k regex:field_cluster_kernel -s 1 -c 1 -o gpurun_out/r01_k3_cluster -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_k3.log 2>&1
# K7 (full-Ewald polarization): real-space cluster kernel and the dipole structure factor
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --workload mof_h2_63k_ewald_full > gpurun_out/ncu_plain_k7.log 2>&1
$NCU -k regex:"induced_real_cluster_kernel|dipole_sf_kernel|induced_finish_kernel" -s 3 -c 3 -o gpurun_out/r01_k7 -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload mof_h2_63k_ewald_full > gpurun_out/ncu_k7.log 2>&1
for f in r01_k4_v8 r01_k1_cluster r01_k3_cluster r01_k7; do
  ncu -i gpurun_out/$f.ncu-rep --page details --csv > gpurun_out/$f.details.csv 2>/dev/null
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.raw.csv 2>/dev/null
done
tail -1 gpurun_out/ncu_plain.log | cut -c1-300; tail -1 gpurun_out/ncu_plain_k7.log | cut -c1-1500
ls -la gpurun_out/*.ncu-rep
