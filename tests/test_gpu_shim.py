"""The drop-in boundary on the GPU: the reference's unmodified host program (parser, MC driver,
Mersenne RNG, writers) linked against mpmc_b200/shim/energy_shim.c + libmpmc_b200.so
(mpmc_b200/shim/_build/, built where /root/reference is present and shipped to the GPU box).

  * single-point energy()/polar() through system_t: shim_driver (oracle/ref_driver.c linked against
    the drop-in) prints what energy() left in system_t / atom_t; it must equal the fixtures
    generated from the unmodified reference (tests/golden/*.npz);
  * fixed-seed MC trajectories (NVT and uVT with insert/remove): `mpmc_b200 <input>` must print
    the same energy.dat rows -- i.e. the same accept/reject sequence -- and the same final
    configuration as the reference's `mpmc` (tests/golden/traj_*.json)."""
import json
import os
import subprocess

import numpy as np
import pytest

from mpmc_b200 import systems
from tests import golden_util as gu

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "mpmc_b200", "shim", "_build")
DRIVER = os.path.join(BUILD, "shim_driver")
PROGRAM = os.path.join(BUILD, "mpmc_b200")

SINGLE_POINT = ["ref_small", "ref_socmof", "ref_socmof_production", "mixed_default", "mixed_tri_default", "mixed_gs_ranked",
                "mixed_palmo", "mixed_precision", "mixed_polar_ewald", "mixed_wolf_full", "mixed_wh", "mixed_sor",
                "mixed_autoreject_fires", "mixed_nopolar_kmax5", "mixed_rd_only", "co2_fluid_375", "h2_fluid_500",
                "ref_test_lenard_jones", "ref_test_coulombic", "mixed_ewald_full", "mixed_ewald_full_precision_palmo_rrms",
                "mixed_tri_ewald_full", "mixed_es_wolf", "mixed_tri_es_wolf_nopolar",
                "mixed_rd_crystal2", "mixed_tri_rd_crystal2_fh2", "ref_pcn61_h2_500"]


def _need(path):
    if not os.path.exists(path):
        pytest.fail(f"{path} missing: run `python -c 'import __graft_entry__ as g; g.build()'` where /root/reference exists")


@pytest.mark.parametrize("name", SINGLE_POINT)
def test_dropin_energy_matches_reference_fixture(name, tmp_path):
    from oracle import oracle  # snapshot parser only (test infrastructure)

    _need(DRIVER)
    s, exp = gu.load(name)
    # fixtures keep molecule_t.mass only (what the hot path reads): put it on the first atom of each
    # molecule so that update_com() recovers the same molecular mass
    first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
    s.meta["atom_mass"] = np.where(first, s.mol_mass, 0.0)
    inp = systems.write_reference_case(s, str(tmp_path))
    r = subprocess.run([DRIVER, os.path.basename(inp)], cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MPMC_SNAPSHOT" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    _, got, _ = oracle.parse_snapshot(r.stdout, name)
    assert got["returned"] == got["total"] or got["total"] >= 1e40 or exp.get("total", 0) >= 1e40
    for k in ("rd", "coulombic", "polar", "total"):
        b = exp[k]
        tol = 1e-10 * abs(b)
        if k in ("coulombic", "total"):
            tol += 1e-13 * max(abs(exp.get("es_real", 0.0)), abs(exp.get("es_self", 0.0))) + 1e-10 * (abs(exp["rd"]) + abs(exp["polar"])) * (k == "total")
        assert abs(got[k] - b) <= tol, f"{name} {k}: drop-in {got[k]!r} vs reference {b!r}"
    assert int(got["polar_iterations"]) == int(exp["polar_iterations"])
    assert int(got["iter_success"]) == int(exp["iter_success"])
    if s.params.polarization and not s.params.rd_only and exp["total"] < 1e40:
        gu.assert_atom_arrays(got, exp, s, rtol=1e-10, label=name)


@pytest.mark.parametrize("host_path", ["cached_sequence", "list_walks", "incremental", "warm_start", "threaded"])
@pytest.mark.parametrize("name", ["traj_small_nvt", "traj_small_uvt", "traj_co2_nvt", "traj_co2_uvt", "traj_co2_npt",
                                  "traj_co2_uvt_cavity", "traj_co2_uvt_insert_input"])
def test_dropin_fixed_seed_trajectory(name, host_path, tmp_path):
    """host_path: the shim's one-pass O(N) duties over the cached pointer sequence (default; every relinking of the
    lists -- uVT insert / remove, rejected moves -- must be detected; MPMC_B200_SHIM_CHECK=1 compares the pass with a real
    list walk after every call), the reference-shaped list walks, or the pass and the scatter with host threads.
    traj_co2_npt: volume moves change the cell, the cutoff, alpha and the k-vectors every step (mc_moves.c:168-232)."""
    _need(PROGRAM)
    env = {"MPMC_B200_HOST_THREADS": "1"}
    if host_path == "cached_sequence":
        env["MPMC_B200_SHIM_CHECK"] = "1"
    if host_path == "list_walks":
        env["MPMC_B200_SHIM_SLOW"] = "1"
    if host_path == "incremental":  # applies to traj_co2_nvt; the polarizable cases must not change
        env["MPMC_B200_INCREMENTAL"] = "1"
    if host_path == "warm_start":  # opt-in warm start: never applied to these fixed-iteration runs, which must not change
        env["MPMC_B200_WARM_START"] = "1"
    if host_path == "threaded":
        env.update({"MPMC_B200_HOST_THREADS": "4", "MPMC_B200_HOST_THREADS_MIN": "0", "MPMC_B200_SHIM_CHECK": "1"})
    gu.run_and_check_trajectory(PROGRAM, name, tmp_path, env)


@pytest.mark.parametrize("name,batch", [("traj_small_nvt", 8), ("traj_small_nvt", 64), ("traj_co2_nvt", 16), ("traj_co2_uvt", 8),
                                        ("traj_co2_npt", 8)])
def test_dropin_resident_steps(name, batch, tmp_path):
    """SURVEY 8f f3 (opt-in MPMC_B200_RESIDENT_STEPS): batches of NVT displacement steps moved, evaluated and decided on the
    device (mpmc_b200_mc_steps), served to the unmodified mc() after checking the trial coordinates bit for bit.  The
    reference's fixed-seed trajectory must come out row for row, and -- the point of running it on the GPU -- the device's
    move arithmetic and its Metropolis decisions must agree with make_move() / mc() so that no batch is cut short (a cut
    batch is still correct, it just falls back to the normal path).  uVT is not eligible and must be untouched."""
    import re

    _need(PROGRAM)
    env = {"MPMC_B200_RESIDENT_STEPS": str(batch), "MPMC_B200_SHIM_CHECK": "1", "MPMC_B200_SHIM_TIMING": "1", "MPMC_B200_HOST_THREADS": "1"}
    r = gu.run_and_check_trajectory(PROGRAM, name, tmp_path, env)
    steps = int(re.findall(r"^numsteps\s+(\d+)", json.load(open(os.path.join(gu.GOLD, name + ".json")))["input"], re.M)[-1])
    m = re.search(r"MPMC_B200_RESIDENT_STEPS: (\d+) batches, (\d+) steps answered from a batch, (\d+) batches cut short", r.stderr)
    if "nvt" in name:
        assert m, r.stderr[-2000:]
        batches, served, dropped = map(int, m.groups())
        assert served == steps and dropped == 0 and batches == -(-steps // batch), (batches, served, dropped, steps)
    elif "npt" in name:  # displacement steps between the volume moves; a batch ends where the driver will draw a volume move
        assert m, r.stderr[-2000:]
        batches, served, dropped = map(int, m.groups())
        assert served > steps // 2 and dropped == 0, (batches, served, dropped, steps)
    else:
        assert m is None or int(m.group(2)) == 0


def test_dropin_energy_no_observables(tmp_path):
    """energy_no_observables() (energy.c:232-318) through the drop-in: same energy, observables_t left alone."""
    from oracle import oracle  # snapshot parser only

    _need(DRIVER)
    s, exp = gu.load("mixed_default")
    first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
    s.meta["atom_mass"] = np.where(first, s.mol_mass, 0.0)
    inp = systems.write_reference_case(s, str(tmp_path))
    r = subprocess.run([DRIVER, os.path.basename(inp), "1", "0", "n"], cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MPMC_SNAPSHOT" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    _, got, _ = oracle.parse_snapshot(r.stdout, "mixed_default")
    assert abs(got["returned"] - exp["total"]) <= 1e-10 * (abs(exp["rd"]) + abs(exp["polar"]) + abs(exp["coulombic"]))
    assert got["total"] == 0.0 and got["rd"] == 0.0 and got["polar"] == 0.0
    gu.assert_atom_arrays(got, exp, s, rtol=1e-10, label="no_observables")


def test_dropin_pcn61_uvt_averages_and_uptake(tmp_path):
    """north_star: "the ensemble averages and uptake of the sample MOF runs must agree".  The shipped PCN-61 + H2 uVT input
    (sample_configs_gpu/3_PCN61/iter.inp, every option of it) on the largest part of the system the reference can hold
    (framework + the first 500 H2: 8,548 atoms), 100 fixed-seed steps with insertions / removals: every energy row, the
    final configuration AND every number of every `OUTPUT:` averages block (N, wt %, excess adsorption ratio, <U>, <U_es>,
    <U_rd>, <U_pol> with its iteration count, qst, acceptance ratios) equal the unmodified reference's
    (tests/golden/traj_pcn61_uvt.json, oracle/make_golden.py traj_pcn61_uvt)."""
    _need(PROGRAM)
    if not os.path.exists(os.path.join(gu.GOLD, "traj_pcn61_uvt.json")):
        pytest.fail("tests/golden/traj_pcn61_uvt.json missing: python oracle/make_golden.py traj_pcn61_uvt")
    gu.run_and_check_trajectory(PROGRAM, "traj_pcn61_uvt", tmp_path, {"MPMC_B200_SHIM_CHECK": "1"})


def _gpu_count():
    import ctypes

    try:
        cuda = ctypes.CDLL("libcudart.so")
    except OSError:
        try:
            import torch

            return torch.cuda.device_count()
        except Exception:
            return 1
    n = ctypes.c_int(0)
    return n.value if cuda.cudaGetDeviceCount(ctypes.byref(n)) == 0 else 1


@pytest.mark.parametrize("name", ["traj_small_nvt", "traj_small_uvt"])
def test_dropin_trajectory_on_a_device_group(name, tmp_path):
    """MPMC_B200_DEVICES=0,1: one MC chain sharded over two GPUs from ONE process (mpmc_b200_create_group) reproduces the
    reference's fixed-seed trajectory row for row, insertions / removals included (the group re-wires its exchange
    buffers whenever the padded sizes change)."""
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs")
    _need(PROGRAM)
    gu.run_and_check_trajectory(PROGRAM, name, tmp_path, {"MPMC_B200_DEVICES": "0,1", "MPMC_B200_SHIM_CHECK": "1"})
