#!/bin/bash
# round 2, GPU call 8 (two GPUs): GS resolver v3 (coefficient tables, branch-free chain); per-rank host threads of a device group
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pins.py -m gpu -q -x > gpurun_out/r02_c8_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c8_tests.log
MPMC_B200_GS_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 2 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c8_bench_gs.json 2> gpurun_out/r02_c8_bench_gs.err
timeout 600 python -m pytest tests/test_gpu_group.py -m gpu -q > gpurun_out/r02_c8_group_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c8_group_tests.log
timeout 300 python -m pytest tests/test_gpu_shim.py -m gpu -q -k "device_group" >> gpurun_out/r02_c8_group_tests.log 2>&1
python - > gpurun_out/r02_c8_dropin_group_timing.txt 2>&1 <<'PY'
import os, subprocess, sys, tempfile
sys.path.insert(0, os.getcwd())
import bench
from mpmc_b200 import systems
s = bench.build_workload("mof_h2_63k")
with tempfile.TemporaryDirectory() as td:
    inp = systems.write_reference_case(s, td)
    for env in ({"MPMC_B200_DEVICES": "0,1"}, {"MPMC_B200_DEVICES": "0,1", "MPMC_B200_GROUP_THREADS": "0"}, {}):
        r = subprocess.run([bench.SHIM_DRIVER, "input", "20", "3", "q"], cwd=td, capture_output=True, text=True,
                           env=dict(os.environ, MPMC_B200_SHIM_TIMING="1", **env))
        print(env, [l for l in r.stdout.splitlines() if l.startswith("timing")], r.stderr[-400:])
PY
tail -3 gpurun_out/r02_c8_tests.log; grep GS_TRACE gpurun_out/r02_c8_bench_gs.err | tail -1; cut -c1-200 gpurun_out/r02_c8_bench_gs.json; tail -4 gpurun_out/r02_c8_group_tests.log; cat gpurun_out/r02_c8_dropin_group_timing.txt
