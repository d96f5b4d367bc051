"""mpmc_b200_mc_steps through the C ABI: a batch of displacement moves applied, evaluated and decided on the
device.  Checked against (a) a numpy restatement of translate() / rotate() (src/mc/mc_moves.c:376-464) -- every numpy
operation rounds once, like the reference's gcc objects and like the _rn intrinsics of mc_step_kernel.cu, so the trial
coordinates must agree BIT FOR BIT; (b) the engine's own mpmc_b200_energy() at those coordinates; (c) the Metropolis rule
replayed on the host; (d) the oracle for the first and the last configuration."""
import ctypes as C

import numpy as np
import pytest

from mpmc_b200 import systems
from mpmc_b200.api import McMove, McResult

pytestmark = pytest.mark.gpu
MAXA, MAXS = 16, 64


def _qmul(a, b):  # quaternion.c:76-86, (x, y, z, w), left to right
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    w = aw * bw - ax * bx - ay * by - az * bz
    x = aw * bx + ax * bw + ay * bz - az * by
    y = aw * by - ax * bz + ay * bw + az * bx
    z = aw * bz + ax * by - ay * bx + az * bw
    return (x, y, z, w)


def _host_move(pos, mass, trans, quat):
    """translate() + rotate() of one molecule, python floats (IEEE double, one rounding per operation)"""
    pos = [[float(v) for v in r] for r in pos]
    m, com = 0.0, [0.0, 0.0, 0.0]
    for i, r in enumerate(pos):
        m += float(mass[i])
        for p in range(3):
            com[p] += float(mass[i]) * r[p]
    com = [c / m for c in com]
    com = [com[p] + float(trans[p]) for p in range(3)]
    pos = [[r[p] + float(trans[p]) for p in range(3)] for r in pos]
    q = tuple(float(v) for v in quat)
    qc = (-q[0], -q[1], -q[2], q[3])
    out = []
    for r in pos:
        pv = (r[0] - com[0], r[1] - com[1], r[2] - com[2], 0.0)
        a = _qmul(q, _qmul(pv, qc))
        out.append([a[0] + com[0], a[1] + com[1], a[2] + com[2]])
    return np.array(out)


@pytest.mark.parametrize("which", ["h2_fluid_polar", "co2_fluid_nonpolar", "mof_h2_palmo"])
def test_mc_steps_match_host_replay(engine, which):
    rng = np.random.default_rng(20261105)
    if which == "h2_fluid_polar":
        s = systems.h2_fluid(1000)
    elif which == "co2_fluid_nonpolar":
        s = systems.co2_fluid(500)
    else:
        s = systems.mof_h2(cells=(3, 3, 1))
        s.params.polar_palmo = 1
    s.params.temperature = s.params.temperature if s.params.temperature > 0 else 77.0
    T = 77.0
    polar = bool(s.params.polarization and not s.params.rd_only)
    engine.set_system(s)
    e0 = engine.energy()
    # movable molecules: contiguous runs of mol_id, not frozen
    starts = np.flatnonzero(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]])
    ends = np.r_[starts[1:], s.n]
    movable = [(int(a), int(b)) for a, b in zip(starts, ends) if not s.frozen[a] and b - a <= MAXA]
    B = 12
    moves = (McMove * B)()
    atom_mass = np.asarray(s.meta.get("atom_mass", np.ones(s.n)), dtype=np.float64)
    atom_mass = np.where(atom_mass > 0, atom_mass, 1.0)
    picks = []
    for j in range(B):
        a, b = movable[int(rng.integers(len(movable)))] if j != 5 else picks[4]  # step 5 moves step 4's molecule again
        picks.append((a, b))
        mv = moves[j]
        mv.first, mv.count, mv.explicit_positions = a, b - a, 0
        mv.u_accept = float(rng.random())
        scale = 0.02 if j % 3 else 1.5  # small moves are mostly accepted, big ones mostly rejected
        for p in range(3):
            mv.trans[p] = float((rng.random() - 0.5) * scale)
        ax = rng.random(3) - 0.5
        ang = float(rng.random() * 2 * np.pi * (0.05 if j % 3 else 1.0))
        ax = ax / np.sqrt((ax * ax).sum())
        for p in range(3):
            mv.quat[p] = float(ax[p] * np.sin(ang / 2))
        mv.quat[3] = float(np.cos(ang / 2))
        for i in range(b - a):
            mv.mass[i] = float(atom_mass[a + i])
    res = (McResult * B)()
    lib = engine.lib
    rc = lib.mpmc_b200_mc_steps(engine._ctx, B, moves, C.c_double(T), C.c_double(e0.total), res)
    assert rc == 0, lib.mpmc_b200_last_error().decode()
    after = engine.energy()  # the engine is left at the configuration after the last decision
    arrays = []
    if polar:
        for j in range(B):
            ptr = [C.POINTER(C.c_double)() for _ in range(6)]
            assert lib.mpmc_b200_mc_step_arrays_view(engine._ctx, j, *[C.byref(q) for q in ptr]) == 0
            arrays.append(np.ctypeslib.as_array(ptr[0], shape=(3 * s.n,)).copy())

    # host replay
    pos = s.pos.copy()
    e_cur = e0.total
    n_acc = 0
    for j in range(B):
        a, b = picks[j]
        mv = moves[j]
        want = _host_move(pos[a:b], [mv.mass[i] for i in range(b - a)], list(mv.trans), list(mv.quat))
        got_new = np.array(res[j].new_pos[:3 * (b - a)]).reshape(-1, 3)
        got_old = np.array(res[j].old_pos[:3 * (b - a)]).reshape(-1, 3)
        assert np.array_equal(got_old, pos[a:b]), f"step {j}: the device moved from other coordinates"
        assert np.array_equal(got_new, want), f"step {j}: trial coordinates differ from translate()/rotate() by {np.abs(got_new - want).max():.3e}"
        trial = pos.copy()
        trial[a:b] = want
        engine.set_positions(0, trial)
        e = engine.energy()
        for k in ("rd", "coulombic", "polar", "total"):
            x, y = getattr(res[j].e, k), getattr(e, k)
            assert abs(x - y) <= 1e-12 * max(abs(y), 1.0), f"step {j} {k}: batch {x!r} vs energy() {y!r}"
        assert res[j].e.polar_iterations == e.polar_iterations
        if polar:
            mu = engine.atom_arrays()["mu"].reshape(-1)
            assert np.abs(arrays[j] - mu).max() <= 1e-12 * np.abs(mu).max()
        bf = np.exp(-(res[j].e.total - e_cur) / T) if np.isfinite(res[j].e.total) else 0.0
        accept = bool(mv.u_accept < bf) and res[j].e.iter_success == 0
        if abs(mv.u_accept - bf) > 1e-12 * max(bf, 1e-300):  # (not within an ulp of the threshold)
            assert bool(res[j].accepted) == accept, f"step {j}: decision"
        if res[j].accepted:
            pos = trial
            e_cur = res[j].e.total
            n_acc += 1
    assert 0 < n_acc < B, "the test should see both accepted and rejected steps"
    # a plain energy() right after the batch (above) must be the energy of the replayed final configuration
    engine.set_positions(0, pos)
    again = engine.energy()
    assert abs(after.total - again.total) <= 1e-12 * abs(again.total)
    from oracle import oracle

    s2 = s.copy()
    s2.pos = pos
    exp, _, _ = oracle.energy(s2, nthreads=0)
    assert abs(again.total - exp.total) <= 1e-10 * (abs(exp.rd) + abs(exp.polar) + abs(exp.coulombic))


def test_mc_steps_refuses_what_it_does_not_cover(engine):
    s = systems.h2_fluid(500)
    s.params.polar_max_iter = 0
    s.params.polar_precision = 1e-6
    engine.set_system(s)
    moves = (McMove * 1)()
    moves[0].first, moves[0].count = 0, 2
    res = (McResult * 1)()
    assert engine.lib.mpmc_b200_mc_steps(engine._ctx, 1, moves, C.c_double(77.0), C.c_double(0.0), res) == 3
    assert b"fixed-iteration" in engine.lib.mpmc_b200_last_error()
    s.params.polar_max_iter, s.params.polar_precision = 10, 0.0
    engine.set_system(s)
    assert engine.lib.mpmc_b200_mc_steps(engine._ctx, MAXS + 1, moves, C.c_double(77.0), C.c_double(0.0), res) == 1
    moves[0].count = MAXA + 1
    assert engine.lib.mpmc_b200_mc_steps(engine._ctx, 1, moves, C.c_double(77.0), C.c_double(0.0), res) == 1
