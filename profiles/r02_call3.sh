#!/bin/bash
# round 2, GPU call 3 (one GPU): persistent Gauss-Seidel sweep, full suite, shard-size probe of K4, ensembles workload
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "gauss or gs" > gpurun_out/r02_c3_gs_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c3_gs_tests.log
timeout 300 python bench.py --steps 5 --warmup 3 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c3_bench_gs.json 2> gpurun_out/r02_c3_bench_gs.err
MPMC_B200_GS=blocks timeout 300 python bench.py --steps 5 --warmup 3 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c3_bench_gs_blocks.json 2> gpurun_out/r02_c3_bench_gs_blocks.err
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r02_c3_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_c3_gputests.log
timeout 300 python profiles/k4_shard_probe.py 8 4 > gpurun_out/r02_c3_shard_probe.txt 2>&1
timeout 300 python bench.py --gpus 1 --steps 5 --warmup 3 --workload ensembles_h2_fluid_16000 --no-cpu-baseline > gpurun_out/r02_c3_bench_ens1.json 2> gpurun_out/r02_c3_bench_ens1.err
tail -4 gpurun_out/r02_c3_gs_tests.log; tail -4 gpurun_out/r02_c3_gputests.log; cut -c1-300 gpurun_out/r02_c3_bench_gs.json; cat gpurun_out/r02_c3_shard_probe.txt | tail -14
