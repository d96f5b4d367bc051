#!/bin/bash
# round 2, GPU call 33 (one GPU): ncu --set full of K1 after the arithmetic work: walk form (co2_fluid_12k, cutoff L/2) and
# compact form (mof_h2_63k, 8 A cutoff); K3 at h2_fluid_16000.  Each after its plain run exited 0.
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-mc-dropin --workload co2_fluid_12k > gpurun_out/r02_c33_plain_co2.log 2>&1 || exit 1
$NCU -k regex:pair_cluster_kernel -s 2 -c 1 -o gpurun_out/r02_k1_walk_co2 -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-mc-dropin --workload co2_fluid_12k > gpurun_out/r02_c33_ncu1.log 2>&1
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c33_plain_mof.log 2>&1 || exit 1
$NCU -k regex:pair_cluster_kernel -s 2 -c 1 -o gpurun_out/r02_k1_compact_mof -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c33_ncu2.log 2>&1
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-mc-dropin --workload h2_fluid_16000 > gpurun_out/r02_c33_plain_h2.log 2>&1 || exit 1
MPMC_B200_OVERLAP=0 $NCU -k regex:"pair_cluster_kernel|field_cluster_kernel" -s 4 -c 2 -o gpurun_out/r02_k1k3_h2_16000_final -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-mc-dropin --workload h2_fluid_16000 > gpurun_out/r02_c33_ncu3.log 2>&1
for f in r02_k1_walk_co2 r02_k1_compact_mof r02_k1k3_h2_16000_final; do
  ncu -i gpurun_out/$f.ncu-rep --page details --csv > gpurun_out/${f}_ncu_details.csv 2>/dev/null
done
ls -la gpurun_out/*.ncu-rep | tail -4
