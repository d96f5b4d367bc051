"""Opt-in warm start of the tolerance-mode dipole solve (include/mpmc_b200.h: mpmc_b200_set_warm_start).

Two contexts on the same GPU follow the same small moves: one starts every solve from init_dipoles()
(thole_iterative.c:13-25), the other from the previous evaluation's converged dipoles.  Both must stop at the same
tolerance -- dipoles within a few times polar_precision of each other -- and the warm one in fewer passes."""
import numpy as np
import pytest

from mpmc_b200 import Engine, systems
from mpmc_b200.api import DEBYE2SKA

pytestmark = pytest.mark.gpu


def _system(kind, solver):
    # the MOF supercell (frozen, strongly polarized framework + mobile H2) is where the cold start is far from the
    # solution everywhere while a single-molecule move disturbs it locally; in the pure fluid the two starts are close
    s = systems.mof_h2(cells=(3, 3, 2)) if kind == "mof" else systems.h2_fluid(4000)
    p = s.params
    p.polar_max_iter, p.polar_precision = 0, 1e-6
    if solver == "gs":
        p.polar_gs = 1
    if solver == "jacobi_palmo_rrms":
        p.polar_palmo, p.polar_rrms = 1, 1
    return s


@pytest.mark.parametrize("kind,solver", [("mof", "jacobi"), ("mof", "gs"), ("mof", "jacobi_palmo_rrms"), ("h2", "jacobi")])
def test_warm_start_same_tolerance_fewer_passes(kind, solver):
    s = _system(kind, solver)
    tol = s.params.polar_precision * DEBYE2SKA  # are_we_done_yet(): thole_iterative.c:108-111
    warm, cold = Engine(0), Engine(0)
    try:
        warm.set_warm_start(True)
        warm.set_system(s)
        cold.set_system(s)
        w, c = warm.energy(), cold.energy()
        assert w.total == c.total and w.polar_iterations == c.polar_iterations  # nothing to start from yet
        rng = np.random.default_rng(7)
        pos = s.pos.copy()
        mols = [m for m in np.unique(s.mol_id) if not s.frozen[s.mol_id == m].all()]
        passes_w = passes_c = 0
        for step in range(4):
            a = np.flatnonzero(s.mol_id == mols[rng.integers(len(mols))])
            pos[a] += rng.normal(size=3) * 0.15
            lo, hi = int(a.min()), int(a.max())
            warm.set_positions(lo, pos[lo:hi + 1])
            cold.set_positions(lo, pos[lo:hi + 1])
            w, c = warm.energy(), cold.energy()
            assert w.iter_success == 0 and c.iter_success == 0
            assert 0 < w.polar_iterations <= c.polar_iterations, f"step {step}: {w.polar_iterations} vs {c.polar_iterations} passes"
            passes_w += w.polar_iterations
            passes_c += c.polar_iterations
            aw, ac = warm.atom_arrays(), cold.atom_arrays()
            dmu = np.abs(aw["mu"] - ac["mu"]).max()
            assert dmu <= 20 * tol, f"step {step}: dipoles differ by {dmu:.3e} (tolerance {tol:.3e})"
            bound = 0.5 * 20 * tol * np.abs(ac["ef_static"]).sum() * (4 if s.params.polar_palmo else 1)
            assert abs(w.polar - c.polar) <= bound, f"step {step}: U_pol {w.polar!r} vs {c.polar!r}"
            for k in ("rd", "es_real", "es_recip", "es_self"):  # nothing else may change
                assert getattr(w, k) == getattr(c, k)
        if kind == "mof":
            assert passes_w < passes_c, f"warm start saved nothing: {passes_w} vs {passes_c} passes over 4 moves"
        # a new atom set has nothing to start from
        warm.set_system(s)
        cold.set_system(s)
        assert warm.energy().polar_iterations == cold.energy().polar_iterations
    finally:
        warm.close()
        cold.close()


def test_warm_start_leaves_fixed_iteration_runs_alone():
    """polar_max_iter runs depend on the starting point: they keep init_dipoles() and stay bit-identical."""
    s = systems.h2_fluid(1000)
    assert s.params.polar_precision == 0.0 and s.params.polar_max_iter > 0
    warm, cold = Engine(0), Engine(0)
    try:
        warm.set_warm_start(True)
        warm.set_system(s)
        cold.set_system(s)
        warm.energy(), cold.energy()
        pos = s.pos.copy()
        pos[:5] += 0.1
        warm.set_positions(0, pos[:5])
        cold.set_positions(0, pos[:5])
        w, c = warm.energy(), cold.energy()
        assert w.total == c.total and w.polar == c.polar and w.polar_iterations == c.polar_iterations
        assert np.array_equal(warm.atom_arrays()["mu"], cold.atom_arrays()["mu"])
    finally:
        warm.close()
        cold.close()
