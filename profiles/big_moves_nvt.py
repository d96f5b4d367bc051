"""NVT through the drop-in with small and with large displacement moves (move_factor scales the cutoff, src/mc/mc_moves.c:380-382;
the reference's default is 1.0).  A molecule that moves further than 1.5 A from where it was when the spatial orders were made
asks for new orders: how much of a step that costs.  usage (GPU box): python profiles/big_moves_nvt.py [workload] [steps]"""
import os
import re
import subprocess
import sys
import tempfile

sys.path.insert(0, os.getcwd())
import numpy as np

import bench
from mpmc_b200 import systems

if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "mof_h2_21k"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
    s = bench.build_workload(name)
    if "atom_mass" not in s.meta:
        s.meta["atom_mass"] = np.where(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]], s.mol_mass, 0.0)
    for mf in ("0.01", "0.3", "1.0"):
        with tempfile.TemporaryDirectory(prefix="mpmc_mv_") as td:
            ex = f"numsteps {steps}\ncorrtime {steps}\npreset_seeds 12345\nmove_factor {mf}\nrot_factor 1.0\n"
            systems.write_reference_case(s, td, ensemble="nvt", extra=ex)
            env = dict(os.environ, MPMC_B200_SHIM_TIMING="1")
            r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
        rate = re.search(r"([\d.]+) energy\(\) calls per second", r.stderr)
        acc = [l for l in r.stdout.splitlines() if "acceptance" in l.lower() or "ar =" in l.lower()]
        print(f"{name} NVT move_factor {mf}: MC steps/s {rate.group(1) if rate else None} rc={r.returncode} {acc[-1][:100] if acc else ''}", flush=True)
