#!/bin/bash
# round 2, GPU call 5 (one GPU): list-based K4 reduce, GS resolver v2
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pins.py -m gpu -q -x > gpurun_out/r02_c5_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c5_tests.log
MPMC_B200_GS_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 2 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c5_bench_gs.json 2> gpurun_out/r02_c5_bench_gs.err
timeout 300 python profiles/k4_shard_probe.py 8 4 > gpurun_out/r02_c5_shard_probe.txt 2>&1
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c5_bench.json 2> gpurun_out/r02_c5_bench.err
tail -3 gpurun_out/r02_c5_tests.log; grep GS_TRACE gpurun_out/r02_c5_bench_gs.err | tail -2; cut -c1-200 gpurun_out/r02_c5_bench_gs.json; cat gpurun_out/r02_c5_shard_probe.txt; cut -c1-200 gpurun_out/r02_c5_bench.json; tail -3 gpurun_out/r02_c5_bench.err
