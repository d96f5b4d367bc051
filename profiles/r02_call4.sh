#!/bin/bash
# round 2, GPU call 4 (one GPU): resolver trace of the persistent GS sweep; K4 at shard size after round-robin items; launch list
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pins.py -m gpu -q -x > gpurun_out/r02_c4_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c4_tests.log
MPMC_B200_GS_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 2 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c4_bench_gs.json 2> gpurun_out/r02_c4_bench_gs.err
timeout 300 python profiles/k4_shard_probe.py 8 4 > gpurun_out/r02_c4_shard_probe.txt 2>&1
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c4_bench.json 2> gpurun_out/r02_c4_bench.err
timeout 200 python profiles/k4_shard_probe.py 8 --one > gpurun_out/r02_c4_probe_one.txt 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 120 --csv --log-file gpurun_out/r02_c4_launches_shard8.csv python profiles/k4_shard_probe.py 8 --one > gpurun_out/r02_c4_ncu.log 2>&1
tail -3 gpurun_out/r02_c4_tests.log; grep GS_TRACE gpurun_out/r02_c4_bench_gs.err | tail -3; cut -c1-200 gpurun_out/r02_c4_bench_gs.json; cat gpurun_out/r02_c4_shard_probe.txt; cut -c1-200 gpurun_out/r02_c4_bench.json
