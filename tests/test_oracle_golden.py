"""Pins the oracle (oracle/mpmc_oracle.c) to the reference: every fixture generated from the
unmodified reference build (tests/golden/*.npz, oracle/make_golden.py) and the reference's own
known-answer values.  CPU only."""
import numpy as np
import pytest

from oracle import oracle
from tests import golden_util as gu

CASES = gu.case_names()


def test_fixtures_present():
    assert len(CASES) >= 40
    for must in ("ref_small", "ref_socmof", "mixed_default", "mixed_gs_ranked_palmo_wolf", "mixed_tri_polar_ewald"):
        assert must in CASES


@pytest.mark.parametrize("name", [c for c in CASES if not c.startswith("ref_socmof")])
@pytest.mark.parametrize("nthreads", [1, 4])
def test_oracle_matches_reference(name, nthreads):
    s, exp = gu.load(name)
    e, c, arr = oracle.energy(s, nthreads=nthreads)
    gu.assert_energies(e.to_dict(), exp, rtol=1e-12, label=name)
    if s.params.polarization and not exp["total"] >= 1e40:
        gu.assert_atom_arrays(arr, exp, s, rtol=1e-12, label=name)


@pytest.mark.parametrize("name", ["ref_socmof", "ref_socmof_production"])
def test_oracle_matches_reference_socmof(name):
    s, exp = gu.load(name)
    e, c, arr = oracle.energy(s, nthreads=0)
    gu.assert_energies(e.to_dict(), exp, rtol=1e-12, label=name)
    gu.assert_atom_arrays(arr, exp, s, rtol=1e-12, label=name)


def test_reference_published_goldens():
    """Step-0 rows checked into the reference: sample_configs_gpu/cuda_pol.small/noncuda_control/
    small.energy.dat:2 and sample_configs_gpu/cuda_pol/noncuda_control/socMOF+BSSP.energy.dat:2
    (6 decimals)."""
    pub = {
        "ref_small": (-22.738394, 1.993547, -24.673480, -0.058461),
        "ref_socmof": (-129043.736570, -8861.486645, -93916.094138, -26266.155787),
    }
    for name, (tot, es, rd, pol) in pub.items():
        s, exp = gu.load(name)
        e, _, _ = oracle.energy(s, nthreads=0)
        for got, want in ((e.total, tot), (e.coulombic, es), (e.rd, rd), (e.polar, pol)):
            assert abs(got - want) < 6e-7 * max(1.0, abs(want) * 1e-6 * 1e6 / max(abs(want), 1.0)) + 5.1e-7, (name, got, want)
        for got, want in ((exp["total"], tot), (exp["coulombic"], es), (exp["rd"], rd), (exp["polar"], pol)):
            assert abs(got - want) < 5.1e-7


def test_reference_known_answer_tests():
    """mpmc_testing/tests/{lenard_jones,coulombic,polar,zero_energy,infinite_energy}_test.txt."""
    s, _ = gu.load("ref_test_lenard_jones")
    e, _, _ = oracle.energy(s, 1)
    assert f"{e.rd:.5f}" == "-1.78212"  # precision exact
    s, _ = gu.load("ref_test_coulombic")
    e, _, _ = oracle.energy(s, 1)
    assert abs(e.coulombic - 19295.2415) <= 0.015
    s, _ = gu.load("ref_test_zero_energy")
    e, _, _ = oracle.energy(s, 1)
    assert e.total <= 0.0 and abs(e.total) < 1e-6
    s, _ = gu.load("ref_test_infinite_energy")
    e, _, _ = oracle.energy(s, 1)
    assert e.rd > 1e10
    # polar_test.txt: +1 e charge 5 A from Ar (alpha 1.6392212), damping off.  The reference test runs the
    # matrix-inversion path; with damping off the only pair is es_excluded (Ar has no charge), so T = 0
    # (thole_matrix.c:92-97) and the iterative path must give the same -1/2 alpha E^2.
    from mpmc_b200.api import System, default_params

    p = default_params(np.eye(3) * 1.0e6)
    p.polarization, p.polar_iterative, p.damp_type, p.rd_lrc = 1, 1, 0, 0
    sysm = System(p, [[49995.0, 50000.0, 50000.0], [50000.0, 50000.0, 50000.0]], [408.7816, 0.0], [0.0, 1.6392212],
                  [0, 0], [0, 0], [1.0, 1.0], [0, 1], [0, 0])
    e, _, _ = oracle.energy(sysm, 1)
    assert abs(e.polar - (-219.13236)) <= 0.002
