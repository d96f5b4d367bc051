"""A uVT-like walk through the C ABI: random removals and insertions of whole H2 molecules (set_atoms every time), energy() after
each; prints the mean set_atoms and energy() time per window of edits, with the spatial orders carried over (default) and rebuilt
every time (MPMC_B200_REUSE_ORDER=0).  usage (GPU box): python profiles/uvt_edit_walk.py [workload] [edits]"""
import os
import sys
import time

sys.path.insert(0, os.getcwd())
import numpy as np

import bench
from mpmc_b200 import Engine

if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "mof_h2_21k"
    edits = int(sys.argv[2]) if len(sys.argv) > 2 else 120
    s = bench.build_workload(name)
    rng = np.random.default_rng(5)
    eng = Engine(0)
    eng.set_system(s)
    e0 = eng.energy()
    fields = ("pos", "charge", "polarizability", "epsilon", "sigma", "mol_mass", "mol_id", "frozen")
    starts = np.flatnonzero(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]])
    movable = [int(a) for a in starts if not s.frozen[a]]
    L = np.array([s.params.basis[q][q] for q in range(3)])
    cur = s.copy()
    t_set = t_en = 0.0
    for k in range(1, edits + 1):
        st = np.flatnonzero(np.r_[True, cur.mol_id[1:] != cur.mol_id[:-1]])
        en = np.r_[st[1:], cur.n]
        mv = [i for i in range(len(st)) if not cur.frozen[st[i]]]
        m = mv[int(rng.integers(len(mv)))]
        a, b = int(st[m]), int(en[m])
        nxt = cur.copy()
        if k % 2:  # removal
            keep = np.r_[np.arange(0, a), np.arange(b, cur.n)]
            for f in fields:
                setattr(nxt, f, getattr(cur, f)[keep].copy())
            nxt.mol_id = np.cumsum(np.r_[0, (nxt.mol_id[1:] != nxt.mol_id[:-1]).astype(np.int32)]).astype(np.int32)
        else:  # insertion of a copy of molecule m at a random place, appended at the end of the list
            shift = rng.uniform(-0.5, 0.5, 3) * L
            for f in fields:
                v = getattr(cur, f)
                add = v[a:b].copy()
                if f == "pos":
                    add = add + shift
                if f == "mol_id":
                    add = np.full(b - a, cur.mol_id.max() + 1, dtype=np.int32)
                setattr(nxt, f, np.concatenate([v, add]))
        t0 = time.perf_counter()
        eng.set_system(nxt)
        t1 = time.perf_counter()
        eng.energy()
        t2 = time.perf_counter()
        t_set += t1 - t0
        t_en += t2 - t1
        cur = nxt
        if k % 20 == 0:
            print(f"{name} edits {k - 19:4d}-{k:4d}: set_params+set_atoms {1e3 * t_set / 20:7.3f} ms  energy() {1e3 * t_en / 20:7.3f} ms  n={cur.n}", flush=True)
            t_set = t_en = 0.0
