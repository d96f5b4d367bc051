"""Cost of mpmc_b200_set_atoms (what every accepted uVT insertion / removal costs: repack, spatial orders, uploads) beside one
energy(), through the C ABI.  usage (GPU box): python profiles/set_atoms_cost.py [workloads ...]"""
import os
import sys
import time

sys.path.insert(0, os.getcwd())
import numpy as np

import bench
from mpmc_b200 import Engine

if __name__ == "__main__":
    names = sys.argv[1:] or ["h2_fluid_1000", "mof_h2_7k", "mof_h2_21k", "mof_h2_63k"]
    eng = Engine(0)
    for name in names:
        s = bench.build_workload(name)
        eng.set_system(s)
        eng.energy()
        # a removal and a re-insertion of the last molecule, alternating: the atom count changes every call
        last = int(np.flatnonzero(s.mol_id == s.mol_id[-1])[0])
        small = s.copy()
        for k in ("pos", "charge", "polarizability", "epsilon", "sigma", "mol_mass", "mol_id", "frozen"):
            setattr(small, k, getattr(s, k)[:last].copy())
        t_set, t_en, reps = 0.0, 0.0, 10
        for r in range(reps):
            for sys_ in (small, s):
                t0 = time.perf_counter()
                eng.set_system(sys_)
                t1 = time.perf_counter()
                eng.energy()
                t2 = time.perf_counter()
                t_set += t1 - t0
                t_en += t2 - t1
        print(f"{name:16s} n={s.n:6d}  set_params+set_atoms {1e3 * t_set / (2 * reps):8.3f} ms   energy() after it {1e3 * t_en / (2 * reps):8.3f} ms",
              flush=True)
