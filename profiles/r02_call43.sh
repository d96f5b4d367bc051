#!/bin/bash
# round 2, GPU call 43 (one GPU): host threads of the drop-in's list pass / scatter beyond 8; Gauss-Seidel production with half items
mkdir -p gpurun_out
python - > gpurun_out/r02_c43_dropin_threads.txt 2>&1 <<'PY'
import os, subprocess, sys, tempfile
sys.path.insert(0, os.getcwd())
import bench
from mpmc_b200 import systems
s = bench.build_workload("mof_h2_63k")
with tempfile.TemporaryDirectory(prefix="mpmc_dropin_") as td:
    inp = systems.write_reference_case(s, td)
    for rep in range(2):
        for th in ("8", "12", "16"):
            env = dict(bench.dropin_env(1), MPMC_B200_SHIM_TIMING="1", MPMC_B200_HOST_THREADS=th)
            r = subprocess.run([bench.SHIM_DRIVER, os.path.basename(inp), "30", "5", "q"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
            t = [l for l in r.stdout.splitlines() if l.startswith("timing seconds_per_energy")]
            print(rep, th, t, [l[l.find('ms per call'):] for l in r.stderr.splitlines() if "ms per call" in l], flush=True)
PY
cat gpurun_out/r02_c43_dropin_threads.txt
timeout 300 python bench.py --steps 10 --warmup 3 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c43_bench_gs.json 2> gpurun_out/r02_c43_bench_gs.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c43_bench_gs.json').read().strip().splitlines()[-1])
print('gs production ms_per_step', d['ms_per_step'], {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'))
PY
