#!/bin/bash
# round 2, GPU call 28 (one GPU): resident steps on NPT (GPU trajectory test), and at 63k atoms with a larger ring
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_resident.py tests/test_gpu_shim.py -m gpu -q -x -k "resident or mc_steps" > gpurun_out/r02_c28_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c28_tests.log
tail -4 gpurun_out/r02_c28_tests.log
python - > gpurun_out/r02_c28_resident_63k.txt 2>&1 <<'PY'
import os, re, subprocess, sys, tempfile
sys.path.insert(0, os.getcwd())
import numpy as np
import bench
from mpmc_b200 import systems
s = bench.build_workload("mof_h2_63k")
if "atom_mass" not in s.meta:
    s.meta["atom_mass"] = np.where(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]], s.mol_mass, 0.0)
for B, ring in ((0, 64), (9, 64), (32, 512)):
    with tempfile.TemporaryDirectory(prefix="mpmc_res_") as td:
        steps = 130
        ex = f"numsteps {steps}\ncorrtime {steps}\npreset_seeds 12345\nmove_factor 0.01\nrot_factor 0.01\n"
        systems.write_reference_case(s, td, ensemble="nvt", extra=ex)
        env = dict(os.environ, MPMC_B200_SHIM_TIMING="1", MPMC_B200_RESIDENT_STEPS=str(B), MPMC_B200_RESIDENT_RING_MB=str(ring))
        r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
    rate = re.search(r"([\d.]+) energy\(\) calls per second", r.stderr)
    res = re.search(r"RESIDENT_STEPS: (\d+) batches, (\d+) steps answered from a batch, (\d+) batches cut short", r.stderr)
    print(f"mof_h2_63k NVT B={B} ring={ring} MB: MC steps/s {rate.group(1) if rate else None}  {res.groups() if res else '-'} rc={r.returncode}", flush=True)
PY
cat gpurun_out/r02_c28_resident_63k.txt
