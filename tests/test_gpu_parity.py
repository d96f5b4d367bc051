"""GPU parity tests: the CUDA path, called through the C ABI (libmpmc_b200.so via ctypes), against
(a) fixtures generated from the unmodified reference and (b) the CPU oracle on the same seeded
inputs, plus size-independent properties at BASELINE.json's full sizes."""
import numpy as np
import pytest

from mpmc_b200 import systems
from tests import golden_util as gu

pytestmark = pytest.mark.gpu

CASES = gu.case_names()


@pytest.mark.parametrize("name", CASES)
def test_cuda_matches_reference_fixture(engine, name):
    """Every fixture of tests/golden: generated from the unmodified reference, except oracle_pcn61_full (BASELINE config 3
    at full size, 21,183 atoms: the reference needs ~80 GB), whose expected values come from the oracle."""
    s, exp = gu.load(name)
    engine.set_system(s)
    got = engine.energy()
    gu.assert_energies(got.to_dict(), exp, rtol=1e-10, label=name)
    if s.params.polarization and not s.params.rd_only and not got.autorejected:
        arr = engine.atom_arrays()
        gu.assert_atom_arrays(arr, exp, s, rtol=1e-10, label=name)
    assert engine.counters().kernel_launches > 0


SYM_SMALL = ["mixed_default", "mixed_tri_default", "mixed_palmo", "mixed_sor", "mixed_esor", "mixed_precision", "mixed_rrms",
             "mixed_wolf_full", "mixed_damp_off", "mixed_damp_linear", "h2_fluid_500", "h2_fluid_500_precision_palmo_rrms",
             "ref_socmof"]


@pytest.mark.parametrize("builder,kw,cutoffs", [
    (systems.co2_fluid, dict(n_mol=1500), ("clusters", "tiles")),
    (systems.mof_h2, dict(cells=(3, 3, 1)), ("clusters",)),
])
def test_es_wolf_matches_oracle_midsize(engine, builder, kw, cutoffs, monkeypatch):
    """Option `wolf` (coulombic_wolf(), coulombic.c:269-308) at a size where K1 runs many cluster pairs per warp; both
    forms of K1 (cluster form and the all-pairs tile form behind MPMC_B200_CUTOFF=tiles)."""
    from oracle import oracle

    s = builder(**kw)
    s.params.es_wolf = 1
    s.params.feynman_hibbs = 0
    exp, _, oarr = oracle.energy(s, nthreads=0)
    assert exp.es_recip == 0.0 and exp.es_self == 0.0 and exp.es_real != 0.0
    from mpmc_b200 import Engine

    for form in cutoffs:
        monkeypatch.setenv("MPMC_B200_CUTOFF", form)  # read when the context is created
        eng = engine if form == "clusters" else Engine(0)
        try:
            eng.set_system(s)
            got = eng.energy()
            gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=f"{s.name} es_wolf {form}")
            assert got.es_recip == 0.0 and got.es_self == 0.0
            if s.params.polarization:
                gu.assert_atom_arrays(eng.atom_arrays(), oarr, s, rtol=1e-10, label=s.name)
        finally:
            if eng is not engine:
                eng.close()


@pytest.mark.parametrize("builder,kw,order", [
    (systems.co2_fluid, dict(n_mol=1500), 2),
    (systems.h2_fluid, dict(n_atoms=2000), 3),
])
def test_rd_crystal_matches_oracle_midsize(engine, builder, kw, order):
    """rd_crystal (lj.c:175-216,264-267): LJ image sums over many tile pairs (diagonal and off-diagonal tiles, ragged
    last tile), same-molecule pairs without their home image, the atoms' own images, the wider LRC cutoff."""
    from oracle import oracle

    s = builder(**kw)
    s.params.rd_crystal = 1
    s.params.rd_crystal_order = order
    exp, _, oarr = oracle.energy(s, nthreads=0)
    engine.set_system(s)
    got = engine.energy()
    gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=f"{s.name} rd_crystal {order}")
    assert abs(got.rd_pair - exp.rd_pair) <= 1e-10 * abs(exp.rd_pair)
    if s.params.polarization:
        gu.assert_atom_arrays(engine.atom_arrays(), oarr, s, rtol=1e-10, label=s.name)
    # switching the option off again on the same context gives the plain LJ sum back
    s.params.rd_crystal = 0
    plain, _, _ = oracle.energy(s, nthreads=0)
    engine.set_system(s)
    got = engine.energy()
    gu.assert_energies(got.to_dict(), plain.to_dict(), rtol=1e-10, label=f"{s.name} rd_crystal off again")
    assert abs(exp.rd_pair - plain.rd_pair) > 1e-5 * abs(plain.rd_pair)


def test_rd_crystal_order_limits(engine):
    s = systems.h2_fluid(500)
    s.params.rd_crystal = 1
    for bad in (0, 5):
        s.params.rd_crystal_order = bad
        with pytest.raises(RuntimeError, match="rd.crystal"):
            engine.set_system(s)
    engine.set_system(systems.h2_fluid(500))


def test_es_wolf_with_feynman_hibbs_is_refused(engine):
    """coulombic.c:293-297: the reference dies on FH + wolf; the engine refuses the parameters (no silent fallback)."""
    s = systems.h2_fluid(500)
    s.params.es_wolf = 1
    assert s.params.feynman_hibbs
    with pytest.raises(RuntimeError, match="es_wolf"):
        engine.set_system(s)
    engine.set_system(systems.h2_fluid(500))  # the context stays usable
    assert engine.energy().total != 0.0


@pytest.mark.parametrize("name", SYM_SMALL)
def test_symmetric_contraction_on_small_fixtures(engine, name, monkeypatch):
    """Below 3,072 padded polarizable atoms a single GPU uses the row form of K4 (latency); MPMC_B200_SYM_MIN=0 forces
    the symmetric kernel so that its damping / wolf / triclinic / padding variants stay pinned to the reference fixtures."""
    monkeypatch.setenv("MPMC_B200_SYM_MIN", "0")
    s, exp = gu.load(name)
    engine.set_system(s)
    got = engine.energy()
    gu.assert_energies(got.to_dict(), exp, rtol=1e-10, label=name)
    gu.assert_atom_arrays(engine.atom_arrays(), exp, s, rtol=1e-10, label=name)


@pytest.mark.parametrize("builder,kw", [
    (systems.h2_fluid, dict(n_atoms=4000)),
    (systems.h2_fluid, dict(n_atoms=5500)),  # 3,300 polarizable atoms: just above the size where the symmetric K4 takes over
    (systems.co2_fluid, dict(n_mol=1500)),
    (systems.mof_h2, dict(cells=(3, 3, 1))),
])
def test_cuda_matches_oracle_midsize(engine, builder, kw):
    from oracle import oracle

    s = builder(**kw)
    engine.set_system(s)
    got = engine.energy()
    exp, _, oarr = oracle.energy(s, nthreads=0)
    gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=s.name)
    if s.params.polarization:
        gu.assert_atom_arrays(engine.atom_arrays(), oarr, s, rtol=1e-10, label=s.name)


@pytest.mark.parametrize("builder,kw,par", [
    (systems.h2_fluid, dict(n_atoms=3000), dict(polar_max_iter=5)),
    (systems.mof_h2, dict(cells=(3, 3, 1)), dict(polar_precision=1e-5, polar_max_iter=0, polar_palmo=1, polar_rrms=1)),
])
def test_ewald_full_matches_oracle_midsize(engine, builder, kw, par):
    """Full-Ewald polarization (ewald_full(), polar_ewald.c:375-418) at sizes with many clusters and j slices."""
    from oracle import oracle

    s = builder(**kw)
    s.params.polar_ewald_full = 1
    for k, v in par.items():
        setattr(s.params, k, v)
    engine.set_system(s)
    got = engine.energy()
    exp, _, oarr = oracle.energy(s, nthreads=0)
    gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=s.name)
    gu.assert_atom_arrays(engine.atom_arrays(), oarr, s, rtol=1e-10, label=s.name)


@pytest.mark.parametrize("par", [
    dict(polar_gs=1, polar_max_iter=5),
    dict(polar_gs_ranked=1, polar_max_iter=4, polar_palmo=1, polar_wolf=1, polar_wolf_alpha=0.13),
    dict(polar_gs=1, polar_precision=1e-6, polar_max_iter=0, polar_rrms=1),
])
def test_gauss_seidel_matches_oracle_midsize(engine, par):
    """Delta-form Gauss-Seidel sweeps (symmetric Jacobi part + block pushes) over ~27 blocks of 128 with a ragged last
    block, against the oracle's in-place sweep (thole_iterative.c:38-53)."""
    from oracle import oracle

    s = systems.mof_h2(cells=(3, 3, 1))
    for k, v in par.items():
        setattr(s.params, k, v)
    engine.set_system(s)
    got = engine.energy()
    exp, _, oarr = oracle.energy(s, nthreads=0)
    gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=s.name)
    gu.assert_atom_arrays(engine.atom_arrays(), oarr, s, rtol=1e-10, label=s.name)
    assert got.polar_iterations == exp.polar_iterations


@pytest.mark.parametrize("form", ["blocks", "panels"])
def test_gauss_seidel_launch_forms_agree(engine, form, monkeypatch):
    """The persistent sweep (one cooperative launch: resolver + pushers, the default) against the per-block launches
    (MPMC_B200_GS=blocks) and the panel form (MPMC_B200_GS=panels) of the same Gauss-Seidel sweep, ranked and with palmo,
    on ~135 blocks with a ragged last one: same iteration count, energies and dipoles to rounding; and the default form
    reproduces itself bit for bit (fixed-order sums, no atomics)."""
    from mpmc_b200 import Engine

    s = systems.mof_h2(cells=(9, 3, 3), polar_max_iter=4)
    s.params.polar_gs_ranked, s.params.polar_palmo = 1, 1
    engine.set_system(s)
    a = engine.energy()
    arr = engine.atom_arrays()
    a2 = engine.energy()
    assert a2.polar == a.polar and np.array_equal(engine.atom_arrays()["mu"], arr["mu"])
    monkeypatch.setenv("MPMC_B200_GS", form)  # read when the context is created
    other = Engine(0)
    try:
        other.set_system(s)
        b = other.energy()
        barr = other.atom_arrays()
    finally:
        other.close()
    assert a.polar_iterations == b.polar_iterations
    assert abs(a.polar - b.polar) <= 1e-11 * abs(b.polar)
    for k in ("mu", "ef_induced", "ef_induced_change"):
        scale = np.abs(barr[k]).max()
        assert np.abs(arr[k] - barr[k]).max() <= 1e-11 * scale, k


def test_polar_entry_point(engine):
    s, exp = gu.load("mixed_default")
    engine.set_system(s)
    got = engine.polar()
    assert abs(got.polar - exp["polar"]) <= 1e-10 * abs(exp["polar"])
    assert got.total == got.polar and got.rd == 0.0


def _close(a: dict, b: dict, rtol=1e-12):
    for k, v in a.items():
        if isinstance(v, float):
            scale = max(abs(a["rd"]), abs(a["es_real"]), abs(a["polar"]), 1.0)
            assert abs(v - b[k]) <= rtol * scale, (k, v, b[k])
        else:
            assert v == b[k], (k, v, b[k])


def test_determinism_and_partial_update(engine):
    """Same inputs => bit-identical outputs run to run; moving one molecule through set_positions
    equals a full re-upload of the moved system (to rounding: the spatial order of the polarizable
    list, hence the summation order, is only rebuilt when an atom has drifted more than 1.5 A from
    where it was sorted -- then the order is again the one a fresh upload produces, bit for bit)."""
    s = systems.h2_fluid(2000)
    engine.set_system(s)
    a = engine.energy().to_dict()
    b = engine.energy().to_dict()
    assert a == b
    moved = s.copy()
    moved.pos[35:40] += np.array([0.11, -0.07, 0.05])
    engine.set_positions(35, moved.pos[35:40])
    c = engine.energy().to_dict()
    assert engine.energy().to_dict() == c
    engine.set_system(moved)
    d = engine.energy().to_dict()
    _close(c, d)
    assert c != a
    # a large displacement triggers the re-sort: identical to a fresh upload
    far = s.copy()
    far.pos[35:40] += np.array([7.0, -3.0, 5.5])
    engine.set_positions(35, far.pos[35:40])
    e = engine.energy().to_dict()
    engine.set_system(far)
    assert engine.energy().to_dict() == e
    engine.set_positions(35, s.pos[35:40])
    engine.set_system(s)
    assert engine.energy().to_dict() == a


def test_lattice_translation_invariance_full_size(engine):
    """BASELINE config 4 size (63k atoms, polarizable): translating every atom by a lattice vector,
    or one molecule by a lattice vector, leaves every energy term unchanged (minimum image +
    periodic structure factor), to rounding."""
    s = systems.mof_h2(polar_max_iter=3)
    assert s.n > 60000
    engine.set_system(s)
    e0 = engine.energy()
    shifted = s.copy()
    basis = np.array(s.params.to_dict()["basis"])
    shifted.pos[-5:] += basis[1]  # last H2 molecule moved by one lattice vector
    engine.set_system(shifted)
    e1 = engine.energy()
    for k in ("rd", "polar", "es_recip"):
        a, b = getattr(e0, k), getattr(e1, k)
        assert abs(a - b) <= 1e-9 * abs(a), (k, a, b)
    floor = 1e-12 * abs(e0.es_real)
    assert abs(e0.coulombic - e1.coulombic) <= 1e-9 * abs(e0.coulombic) + floor
    assert e0.polar_iterations == 3 and e0.polar < 0
    c = engine.counters()
    assert c.n_polarizable > 45000 and c.pairs_dipole == float(c.n_polarizable) * (c.n_polarizable - 1)


def test_edge_cases(engine):
    from mpmc_b200.api import System, default_params

    # two atoms, no charges, no polarizability: pure LJ (reference lenard_jones test geometry)
    p = default_params(np.eye(3) * 20.0)
    p.rd_lrc = 0
    s = System(p, [[0, 0, 0], [5, 5, 5]], [0, 0], [0, 0], [128.326802, 128.326802], [3.371914, 3.371914], [39.948, 39.948],
               [0, 1], [0, 0])
    engine.set_system(s)
    e = engine.energy()
    assert f"{e.rd:.5f}" == "-1.78212"  # mpmc_testing/tests/lenard_jones_test.txt
    assert e.coulombic == 0.0 and e.polar == 0.0
    # a single atom: no pairs at all
    s1 = System(p, [[1, 2, 3]], [0.5], [0.0], [10.0], [3.0], [1.0], [0], [0])
    engine.set_system(s1)
    e = engine.energy()
    assert e.rd == 0.0 and np.isfinite(e.total)
    # exactly one full tile and one tile + 1 atom (padding boundaries)
    for n in (640, 645):
        sN = systems.h2_fluid(n)
        from oracle import oracle

        engine.set_system(sN)
        got = engine.energy()
        exp, _, _ = oracle.energy(sN, nthreads=0)
        gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=f"n={n}")


def test_small_edits_of_the_atom_set_keep_parity(engine):
    """uVT through the ABI: after a removal / an insertion of a molecule set_atoms carries the spatial orders over instead of
    sorting afresh (surviving atoms keep their place, new atoms go in behind their nearest survivor; engine.cu: edit_map).
    The orders only re-associate sums: every edited configuration must still meet the oracle at 1e-10, energies and dipoles."""
    from oracle import oracle

    rng = np.random.default_rng(11)
    s = systems.mof_h2(cells=(3, 3, 1))  # 2,391 atoms: above the 1,024-atom threshold of the carry-over
    fields = ("pos", "charge", "polarizability", "epsilon", "sigma", "mol_mass", "mol_id", "frozen")
    engine.set_system(s)
    engine.energy()
    L = np.array([s.params.basis[q][q] for q in range(3)])
    cur = s
    for k in range(6):
        st = np.flatnonzero(np.r_[True, cur.mol_id[1:] != cur.mol_id[:-1]])
        en = np.r_[st[1:], cur.n]
        mv = [i for i in range(len(st)) if not cur.frozen[st[i]]]
        m = mv[int(rng.integers(len(mv)))]
        a, b = int(st[m]), int(en[m])
        nxt = cur.copy()
        if k % 3 != 2:  # two removals ...
            keep = np.r_[np.arange(0, a), np.arange(b, cur.n)]
            for f in fields:
                setattr(nxt, f, getattr(cur, f)[keep].copy())
        else:  # ... then an insertion in the MIDDLE of the list (a copy of molecule m, displaced, right behind it)
            for f in fields:
                v = getattr(cur, f)
                add = v[a:b].copy()
                if f == "pos":
                    add = add + rng.uniform(-0.5, 0.5, 3) * L
                setattr(nxt, f, np.concatenate([v[:b], add, v[b:]]))
            nxt.mol_id[b:b + (b - a)] = -1
        nxt.mol_id = np.cumsum(np.r_[0, (nxt.mol_id[1:] != nxt.mol_id[:-1]).astype(np.int32)]).astype(np.int32)
        engine.set_system(nxt)
        got = engine.energy()
        exp, _, oarr = oracle.energy(nxt, nthreads=0)
        gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=f"edit {k}")
        mu = engine.atom_arrays()["mu"]
        assert np.abs(mu - oarr["mu"]).max() <= 1e-10 * np.abs(oarr["mu"]).max(), f"edit {k}: dipoles"
        cur = nxt


def test_two_edits_in_one_call_carry_the_orders():
    """A uVT step reaches the engine as TWO edits at different places of the list: the driver's restore of a rejected
    insertion / removal arrives together with the next trial (mc_moves.c:790-799 + :567-741).  set_atoms must read that as
    two small edits (engine.cu: two-pointer walk), not as one long changed span, and carry the orders over; a displaced
    molecule in the same call (replacement) as well.  Fresh process with MPMC_B200_REPACK_TIMING, whose stderr line names
    the path taken; every configuration against the oracle."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
from mpmc_b200 import Engine, systems
from oracle import oracle
rng = np.random.default_rng(23)
s = systems.mof_h2(cells=(3, 3, 1))
fields = ("pos", "charge", "polarizability", "epsilon", "sigma", "mol_mass", "mol_id", "frozen")
eng = Engine(0)
eng.set_system(s)
eng.energy()
L = np.array([s.params.basis[q][q] for q in range(3)])
cur = s
for k in range(4):
    st = np.flatnonzero(np.r_[True, cur.mol_id[1:] != cur.mol_id[:-1]])
    en = np.r_[st[1:], cur.n]
    mv = [i for i in range(len(st)) if not cur.frozen[st[i]]]
    pick = sorted(rng.choice(len(mv), size=3, replace=False))
    (a1, b1), (a2, b2), (a3, b3) = [(int(st[mv[i]]), int(en[mv[i]])) for i in pick]
    nxt = cur.copy()
    # molecule 1 removed, molecule 2 displaced (k odd), a displaced copy of molecule 3 inserted behind it
    for f in fields:
        v = getattr(cur, f)
        add = v[a3:b3].copy()
        mid = v[b1:b3].copy()
        if f == "pos":
            add = add + rng.uniform(-0.5, 0.5, 3) * L
            if k %% 2:
                mid[a2 - b1:b2 - b1] += rng.uniform(-0.3, 0.3, 3)
        setattr(nxt, f, np.concatenate([v[:a1], mid, add, v[b3:]]))
    pos_new = b3 - (b1 - a1)
    nxt.mol_id[pos_new:pos_new + (b3 - a3)] = -1
    nxt.mol_id = np.cumsum(np.r_[0, (nxt.mol_id[1:] != nxt.mol_id[:-1]).astype(np.int32)]).astype(np.int32)
    print("EDIT", k, cur.n, "->", nxt.n, file=sys.stderr, flush=True)
    eng.set_system(nxt)
    got = eng.energy()
    exp, _, oarr = oracle.energy(nxt, nthreads=0)
    for key, want in exp.to_dict().items():
        have = got.to_dict()[key]
        if isinstance(want, float) and key not in ("iterations",):
            assert abs(have - want) <= 1e-10 * max(abs(want), 1.0), (k, key, have, want)
    mu = eng.atom_arrays()["mu"]
    assert np.abs(mu - oarr["mu"]).max() <= 1e-10 * np.abs(oarr["mu"]).max(), k
    cur = nxt
eng.close()
print("EDITS_OK")
""" % root
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=900,
                       env=dict(os.environ, MPMC_B200_REPACK_TIMING="1"))
    assert r.returncode == 0 and "EDITS_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
    after = r.stderr.split("EDIT 0")[1]
    lines = [l for l in after.splitlines() if "MPMC_B200_REPACK_TIMING" in l]
    assert len(lines) == 4 and all("orders carried over" in l for l in lines), lines


def test_large_moves_relocate_atoms_in_the_orders(engine):
    """With the reference's default move_factor (1.0: displacements up to the cutoff) nearly every trial move carries a
    molecule further than the drift limit of the spatial orders.  The engine then takes the molecule's atoms out of the orders
    and puts them back behind their nearest neighbour instead of sorting everything again (engine.cu: refresh_subsets).
    Jumps of several Angstrom, a restore and two molecules moved between evaluations: every configuration against the oracle."""
    from oracle import oracle

    rng = np.random.default_rng(17)
    s = systems.mof_h2(cells=(3, 3, 1))  # 2,391 atoms: above the 1,024-atom threshold of the carry-over
    engine.set_system(s)
    engine.energy()
    st = np.flatnonzero(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]])
    en = np.r_[st[1:], s.n]
    mv = [(int(a), int(b)) for a, b in zip(st, en) if not s.frozen[a]]
    cur = s.copy()
    last = None
    for k in range(7):
        nxt = cur.copy()
        if k == 3 and last is not None:  # a reject: the molecule goes back where it was
            a, b, old = last
            nxt.pos[a:b] = old
            moved = [(a, b)]
        else:
            moved = []
            for _ in range(2 if k == 5 else 1):  # k == 5: two molecules between two evaluations
                a, b = mv[int(rng.integers(len(mv)))]
                last = (a, b, cur.pos[a:b].copy())
                nxt.pos[a:b] = cur.pos[a:b] + rng.uniform(-6.0, 6.0, 3)
                moved.append((a, b))
        for a, b in moved:
            engine.set_positions(a, nxt.pos[a:b])
        got = engine.energy()
        exp, _, oarr = oracle.energy(nxt, nthreads=0)
        gu.assert_energies(got.to_dict(), exp.to_dict(), rtol=1e-10, label=f"jump {k}")
        mu = engine.atom_arrays()["mu"]
        assert np.abs(mu - oarr["mu"]).max() <= 1e-10 * np.abs(oarr["mu"]).max(), f"jump {k}: dipoles"
        cur = nxt

