"""Opt-in incremental energy() after small moves (include/mpmc_b200.h: mpmc_b200_set_incremental).

An MC-like random walk (single-molecule translations + rotations, double moves, rejected moves restored bit for bit) is
evaluated by two contexts on the same GPU: one with the incremental path on, one doing the full evaluation every step.
They must agree to rounding at every step, and the incremental context must really have taken the incremental path."""
import numpy as np
import pytest

from mpmc_b200 import Engine, systems

pytestmark = pytest.mark.gpu


def _case(name):
    if name == "co2_ewald":
        return systems.co2_fluid(500)
    if name == "co2_wolf":
        s = systems.co2_fluid(500)
        s.params.es_wolf, s.params.feynman_hibbs = 1, 0
        return s
    if name == "co2_rd_only":
        s = systems.co2_fluid(500)
        s.params.rd_only = 1
        return s
    if name == "mixed_fh4":  # frozen framework atoms, zero-parameter sites, FH4, LRC
        s = systems.mixed_small()
        s.params.polarization = 0
        return s
    if name == "mixed_tri_fh2_wh":
        s = systems.mixed_small(triclinic=True)
        s.params.polarization, s.params.feynman_hibbs_order, s.params.mixing = 0, 2, 1
        return s
    raise KeyError(name)


def _rotate(xyz, rng, angle):
    axis = rng.normal(size=3)
    axis /= np.linalg.norm(axis)
    c, s_ = np.cos(angle), np.sin(angle)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    R = np.eye(3) + s_ * K + (1 - c) * (K @ K)
    com = xyz.mean(axis=0)
    return (xyz - com) @ R.T + com


def _close(a, b, scale, label):
    tol = 1e-10 * abs(b) + 1e-13 * scale
    assert abs(a - b) <= tol, f"{label}: incremental {a!r} vs full {b!r} (|diff| {abs(a - b):.3e} > {tol:.3e})"


@pytest.mark.parametrize("name", ["co2_ewald", "co2_wolf", "co2_rd_only", "mixed_fh4", "mixed_tri_fh2_wh"])
def test_incremental_random_walk_matches_full(name):
    s = _case(name)
    inc, full = Engine(0), Engine(0)
    try:
        inc.set_incremental(True)
        inc.set_system(s)
        full.set_system(s)
        e0, f0 = inc.energy(), full.energy()
        assert not inc.last_eval_incremental()  # the first evaluation is the base
        assert e0.total == f0.total
        scale = max(abs(f0.rd_pair), abs(f0.es_real), abs(f0.es_self), abs(f0.es_recip), 1.0)
        rng = np.random.default_rng(20261104)
        pos = s.pos.copy()
        mols = [np.flatnonzero(s.mol_id == m) for m in np.unique(s.mol_id)]
        mols = [a for a in mols if not s.frozen[a].all()]
        last = None
        n_inc = 0
        for step in range(80):
            kind = rng.random()
            if kind < 0.2 and last is not None:  # a rejected move: the old coordinates come back bit for bit
                for a, old in last:
                    pos[a] = old
                last = None
            else:
                last = []
                for _ in range(2 if kind > 0.85 else 1):
                    a = mols[rng.integers(len(mols))]
                    last.append((a, pos[a].copy()))
                    pos[a] = _rotate(pos[a], rng, rng.normal() * 0.3) + rng.normal(size=3) * 0.25
            touched = np.concatenate([a for a, _ in last]) if last else np.arange(s.n)
            for eng in (inc, full):
                if step % 3 == 0:  # the caller may hand over everything ...
                    eng.set_positions(0, pos)
                else:              # ... or the range that covers the change, as the drop-in shim does
                    lo, hi = int(touched.min()), int(touched.max())
                    eng.set_positions(lo, pos[lo:hi + 1])
            got, exp = inc.energy(), full.energy()
            n_inc += inc.last_eval_incremental()
            assert not full.last_eval_incremental()
            for k in ("rd", "rd_pair", "es_real", "es_recip", "es_self", "coulombic", "total"):
                _close(getattr(got, k), getattr(exp, k), scale, f"{name} step {step} {k}")
            ci, cf = inc.counters(), full.counters()
            assert ci.pairs_lj == cf.pairs_lj and ci.pairs_es == cf.pairs_es, f"{name} step {step}: pair counts"
        assert n_inc == 80, f"{name}: only {n_inc} of 80 evaluations took the incremental path"
        # new parameters or a new atom set: the next evaluation is a full one again
        inc.set_system(s)
        again = inc.energy()
        assert not inc.last_eval_incremental() and again.total == f0.total
    finally:
        inc.close()
        full.close()


def test_incremental_keeps_out_of_polarizable_and_autoreject_runs():
    """Polarization (every dipole changes), cavity_autoreject_absolute (a minimum over all pairs) and rd_crystal keep the
    full evaluation even with the switch on -- bit-identical to a context without it."""
    cases = []
    cases.append(systems.h2_fluid(500))
    s = systems.co2_fluid(125)
    s.params.cavity_autoreject_absolute, s.params.cavity_autoreject_scale = 1, 0.5
    cases.append(s)
    s = systems.co2_fluid(125)
    s.params.rd_crystal, s.params.rd_crystal_order = 1, 2
    cases.append(s)
    for s in cases:
        inc, full = Engine(0), Engine(0)
        try:
            inc.set_incremental(True)
            inc.set_system(s)
            full.set_system(s)
            inc.energy(), full.energy()
            pos = s.pos.copy()
            a = np.flatnonzero(s.mol_id == s.mol_id[-1])
            pos[a] += 0.2
            inc.set_positions(0, pos)
            full.set_positions(0, pos)
            got, exp = inc.energy(), full.energy()
            assert not inc.last_eval_incremental()
            assert got.total == exp.total and got.rd == exp.rd and got.coulombic == exp.coulombic
        finally:
            inc.close()
            full.close()


def test_incremental_falls_back_after_a_large_move():
    """More than 128 moved atoms (a volume move, a replayed frame): full evaluation, then incremental again."""
    s = systems.co2_fluid(500)
    inc, full = Engine(0), Engine(0)
    try:
        inc.set_incremental(True)
        inc.set_system(s)
        full.set_system(s)
        inc.energy(), full.energy()
        pos = s.pos + 0.01
        inc.set_positions(0, pos)
        full.set_positions(0, pos)
        got, exp = inc.energy(), full.energy()
        assert not inc.last_eval_incremental() and got.total == exp.total
        pos[:3] += 0.1
        inc.set_positions(0, pos[:3])
        full.set_positions(0, pos[:3])
        got, exp = inc.energy(), full.energy()
        assert inc.last_eval_incremental()
        assert abs(got.total - exp.total) <= 1e-10 * abs(exp.total) + 1e-13 * abs(exp.es_self)
    finally:
        inc.close()
        full.close()
