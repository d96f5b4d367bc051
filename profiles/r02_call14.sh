#!/bin/bash
# round 2, GPU call 14 (one GPU): GS Jacobi part on the spatial image; new uVT trajectories on the GPU; ncu of K3 / K1 on a fluid
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02_c14_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c14_tests.log
MPMC_B200_GS_TRACE=1 timeout 200 python bench.py --steps 5 --warmup 3 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c14_bench_gs.json 2> gpurun_out/r02_c14_bench_gs.err
MPMC_B200_GS_JACOBI=order timeout 200 python bench.py --steps 5 --warmup 3 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c14_bench_gs_order.json 2> gpurun_out/r02_c14_bench_gs_order.err
timeout 200 python profiles/k4_probe.py h2_fluid_16000 > gpurun_out/r02_c14_probe.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"field_cluster_kernel|pair_cluster_kernel" -s 2 -c 2 -o gpurun_out/r02_k1k3_fluid16k -f python profiles/k4_probe.py h2_fluid_16000 > gpurun_out/r02_c14_ncu.log 2>&1
tail -4 gpurun_out/r02_c14_tests.log; cut -c60-200 gpurun_out/r02_c14_bench_gs.json; cut -c60-200 gpurun_out/r02_c14_bench_gs_order.json; grep GS_TRACE gpurun_out/r02_c14_bench_gs.err | tail -1; tail -2 gpurun_out/r02_c14_ncu.log
