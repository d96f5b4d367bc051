#!/bin/bash
# round 2, GPU call 1 (one GPU): GPU test suite, bench line, drop-in host-time breakdown, executed-flop capture of K4
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02_c1_smi.txt 2>&1
nproc > gpurun_out/r02_c1_nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_c1_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_c1_gputests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_c1_bench.json 2> gpurun_out/r02_c1_bench.err; echo "bench rc=$?" >> gpurun_out/r02_c1_bench.err
timeout 300 python bench.py --steps 10 --warmup 3 --workload co2_fluid_12k --no-cpu-baseline > gpurun_out/r02_c1_bench_co2.json 2> gpurun_out/r02_c1_bench_co2.err
timeout 300 python bench.py --steps 10 --warmup 3 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c1_bench_gs.json 2> gpurun_out/r02_c1_bench_gs.err
# where a drop-in energy() call spends its host time, 1 and 8 host threads
python - > gpurun_out/r02_c1_dropin_timing.txt 2>&1 <<'PY'
import os, subprocess, sys, tempfile
sys.path.insert(0, os.getcwd())
import bench
from mpmc_b200 import systems
s = bench.build_workload("mof_h2_63k")
with tempfile.TemporaryDirectory() as td:
    inp = systems.write_reference_case(s, td)
    for th in ("1", "4", "8"):
        r = subprocess.run([bench.SHIM_DRIVER, "input", "20", "3", "q"], cwd=td, capture_output=True, text=True,
                           env=dict(os.environ, MPMC_B200_SHIM_TIMING="1", MPMC_B200_HOST_THREADS=th))
        print("threads", th, [l for l in r.stdout.splitlines() if l.startswith("timing")], r.stderr[-600:])
PY
python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c1_k4_probe.log 2>&1 && \
ncu --metrics $(python profiles/capture_k4_executed.py --metrics) --clock-control none -k regex:dipole_sym_kernel -s 10 -c 2 --csv \
    --log-file gpurun_out/r02_k4_executed_mof_h2_63k.csv python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c1_ncu.log 2>&1
echo "ncu rc=$?" >> gpurun_out/r02_c1_ncu.log
tail -3 gpurun_out/r02_c1_gputests.log; cut -c1-600 gpurun_out/r02_c1_bench.json; tail -2 gpurun_out/r02_c1_bench.err
