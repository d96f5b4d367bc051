#!/bin/bash
# round 2, GPU call 19 (one GPU): resident MC steps -- ABI test, drop-in trajectories, MC steps/s A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_resident.py tests/test_gpu_shim.py -m gpu -q -x -k "resident or mc_steps" > gpurun_out/r02_c19_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c19_tests.log
tail -25 gpurun_out/r02_c19_tests.log
timeout 1200 python - > gpurun_out/r02_c19_resident_steps.txt 2>&1 <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import bench
from mpmc_b200 import systems
from tests import golden_util as gu
def fix(s):
    if "atom_mass" not in s.meta:
        s.meta["atom_mass"] = np.where(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]], s.mol_mass, 0.0)
    return s
cases = [("cuda_pol.small 10 atoms", fix(gu.load("ref_small")[0]), 200, 8200),
         ("socMOF+BSSP 1,236 atoms", fix(gu.load("ref_socmof")[0]), 100, 4100),
         ("h2_fluid_1000", fix(systems.h2_fluid(1000)), 100, 4100),
         ("h2_fluid_4000", fix(systems.h2_fluid(4000)), 50, 1050),
         ("co2_fluid_3000 (non-polarizable)", fix(systems.co2_fluid(1000)), 100, 4100),
         ("mof_h2_7k", fix(bench.build_workload("mof_h2_7k")), 20, 420)]
for label, s, a, b in cases:
    for B in ("0", "8", "32", "64"):
        os.environ["MPMC_B200_RESIDENT_STEPS"] = B
        v = bench.dropin_mc_steps_per_s(s, 1, a, b)
        print(f"{label:36s} NVT  MPMC_B200_RESIDENT_STEPS={B:3s} MC steps/s through mpmc_b200 <input>: {v}", flush=True)
PY
cat gpurun_out/r02_c19_resident_steps.txt
