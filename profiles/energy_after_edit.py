"""What an evaluation costs right after an edit of the atom set, against the same evaluation on an unchanged / a displaced system
(C ABI, one GPU).  usage (GPU box): python profiles/energy_after_edit.py [workload] [polar_max_iter]"""
import os
import sys
import time

sys.path.insert(0, os.getcwd())
import numpy as np

import bench
from mpmc_b200 import Engine

if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "mof_h2_7k"
    s = bench.build_workload(name)
    if len(sys.argv) > 2:
        s.params.polar_max_iter = int(sys.argv[2])
    eng = Engine(0)
    eng.set_system(s)
    for _ in range(5):
        eng.energy()
    t0 = time.perf_counter()
    for _ in range(50):
        eng.energy()
    t_same = (time.perf_counter() - t0) / 50
    st = np.flatnonzero(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]])
    en = np.r_[st[1:], s.n]
    mv = [(int(a), int(b)) for a, b in zip(st, en) if not s.frozen[a]]
    rng = np.random.default_rng(3)
    t_pos = t_en = 0.0
    pos = s.pos.copy()
    for k in range(50):
        a, b = mv[int(rng.integers(len(mv)))]
        pos[a:b] += rng.uniform(-0.05, 0.05, 3)
        t0 = time.perf_counter()
        eng.set_positions(a, pos[a:b])
        t1 = time.perf_counter()
        eng.energy()
        t2 = time.perf_counter()
        t_pos += t1 - t0
        t_en += t2 - t1
    print(f"{name} n={s.n} iter={s.params.polar_max_iter}: energy() unchanged {1e3 * t_same:.3f} ms; after set_positions "
          f"{1e3 * t_en / 50:.3f} ms (+ set_positions {1e3 * t_pos / 50:.3f})", flush=True)
    fields = ("pos", "charge", "polarizability", "epsilon", "sigma", "mol_mass", "mol_id", "frozen")
    cur = s.copy()
    cur.pos = pos
    t_set = t_en = t_arr = 0.0
    for k in range(40):
        st = np.flatnonzero(np.r_[True, cur.mol_id[1:] != cur.mol_id[:-1]])
        en = np.r_[st[1:], cur.n]
        mvi = [i for i in range(len(st)) if not cur.frozen[st[i]]]
        m = mvi[int(rng.integers(len(mvi)))]
        a, b = int(st[m]), int(en[m])
        nxt = cur.copy()
        if k % 2:
            keep = np.r_[np.arange(0, a), np.arange(b, cur.n)]
            for f in fields:
                setattr(nxt, f, getattr(cur, f)[keep].copy())
        else:
            L = np.array([s.params.basis[q][q] for q in range(3)])
            for f in fields:
                v = getattr(cur, f)
                add = v[a:b].copy()
                if f == "pos":
                    add = add + rng.uniform(-0.5, 0.5, 3) * L
                setattr(nxt, f, np.concatenate([v[:b], add, v[b:]]))
            nxt.mol_id[b:b + (b - a)] = -1
        nxt.mol_id = np.cumsum(np.r_[0, (nxt.mol_id[1:] != nxt.mol_id[:-1]).astype(np.int32)]).astype(np.int32)
        t0 = time.perf_counter()
        eng.set_system(nxt)
        t1 = time.perf_counter()
        eng.energy()
        t2 = time.perf_counter()
        eng.atom_arrays()
        t3 = time.perf_counter()
        if k >= 4:
            t_set += t1 - t0
            t_en += t2 - t1
            t_arr += t3 - t2
        cur = nxt
    print(f"{name}: after an edit: set_params+set_atoms {1e3 * t_set / 36:.3f} ms, energy() {1e3 * t_en / 36:.3f} ms, "
          f"atom arrays {1e3 * t_arr / 36:.3f} ms", flush=True)
