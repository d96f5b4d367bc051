"""MC steps/s of `mpmc_b200 <input>` (NVT, fixed seed) with and without resident MC steps (MPMC_B200_RESIDENT_STEPS), read from
the shim's own clock (MPMC_B200_SHIM_TIMING: energy() calls per second of wall clock between the 1st and the last call, i.e.
without parsing, CUDA start-up and the first allocations, with everything the unmodified driver does per step).
usage: python profiles/resident_steps_ab.py [batch sizes ...]   (GPU box)"""
import os
import re
import subprocess
import sys
import tempfile

sys.path.insert(0, os.getcwd())
import numpy as np

import bench
from mpmc_b200 import systems
from tests import golden_util as gu


def fix(s):
    if "atom_mass" not in s.meta:
        s.meta["atom_mass"] = np.where(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]], s.mol_mass, 0.0)
    return s


CASES = [("cuda_pol.small (10 atoms)", lambda: fix(gu.load("ref_small")[0]), 20000),
         ("socMOF+BSSP (1,236 atoms)", lambda: fix(gu.load("ref_socmof")[0]), 12000),
         ("h2_fluid_1000", lambda: fix(systems.h2_fluid(1000)), 12000),
         ("h2_fluid_2000", lambda: fix(systems.h2_fluid(2000)), 8000),
         ("h2_fluid_4000", lambda: fix(systems.h2_fluid(4000)), 4000),
         ("co2_fluid 3,000 atoms, non-polarizable", lambda: fix(systems.co2_fluid(1000)), 12000),
         ("co2_fluid_12k (BASELINE config 2)", lambda: fix(bench.build_workload("co2_fluid_12k")), 2049),
         ("mof_h2_7k", lambda: fix(bench.build_workload("mof_h2_7k")), 1537),
         ("mof_h2_21k", lambda: fix(bench.build_workload("mof_h2_21k")), 1025),
         ("mof_h2_63k", lambda: fix(bench.build_workload("mof_h2_63k")), 129)]

if __name__ == "__main__":
    only = os.environ.get("RESIDENT_AB_ONLY")  # substring filter on the case label
    batches = [int(a) for a in sys.argv[1:]] or [0, 8, 32, 64]
    for label, make, steps in CASES:
        if only and only not in label:
            continue
        s = make()
        for B in batches:
            with tempfile.TemporaryDirectory(prefix="mpmc_res_") as td:
                ex = f"numsteps {steps}\ncorrtime {steps}\npreset_seeds 12345\nmove_factor 0.01\nrot_factor 0.01\n"
                systems.write_reference_case(s, td, ensemble="nvt", extra=ex)
                env = dict(os.environ, MPMC_B200_SHIM_TIMING="1", MPMC_B200_RESIDENT_STEPS=str(B),
                           MPMC_B200_RESIDENT_RING_MB=os.environ.get("MPMC_B200_RESIDENT_RING_MB", "512"))
                r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
            rate = re.search(r"([\d.]+) energy\(\) calls per second", r.stderr)
            res = re.search(r"RESIDENT_STEPS: (\d+) batches, (\d+) steps answered from a batch, (\d+) batches cut short", r.stderr)
            whole = re.search(r"whole call ([\d.]+)", r.stderr)
            print(f"{label:40s} B={B:3d}  MC steps/s {float(rate.group(1)) if rate else None!s:>9}  energy() ms/call "
                  f"{whole.group(1) if whole else None}  batches/served/cut {res.groups() if res else '-'}  rc={r.returncode}", flush=True)
