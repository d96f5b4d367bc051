#!/bin/bash
# round 2, GPU call 37 (one GPU): where the drop-in energy(system_t*) call spends its time at 63k atoms (three repeats)
mkdir -p gpurun_out
python - > gpurun_out/r02_c37_dropin_e2e.txt 2>&1 <<'PY'
import os, subprocess, sys, tempfile
sys.path.insert(0, os.getcwd())
import bench
from mpmc_b200 import systems
s = bench.build_workload("mof_h2_63k")
with tempfile.TemporaryDirectory(prefix="mpmc_dropin_") as td:
    inp = systems.write_reference_case(s, td)
    for rep in range(3):
        for env_extra in ({}, {"MPMC_B200_HOST_THREADS": "1"}):
            env = dict(bench.dropin_env(1), MPMC_B200_SHIM_TIMING="1", **env_extra)
            r = subprocess.run([bench.SHIM_DRIVER, os.path.basename(inp), "30", "5", "q"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
            t = [l for l in r.stdout.splitlines() if l.startswith("timing seconds_per_energy")]
            print(rep, env_extra, t, [l for l in r.stderr.splitlines() if "ms per call" in l], flush=True)
PY
cat gpurun_out/r02_c37_dropin_e2e.txt
nproc; cat /proc/cpuinfo | grep "model name" | head -1
