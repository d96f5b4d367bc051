"""The HOST logic of the drop-in shim (mpmc_b200/shim/energy_shim.c) on the CPU: its own objects -- the reference's unmodified
parser, MC driver, Mersenne RNG and writers + energy_shim.c -- linked with tests/hoststub/abi_over_oracle.c, which answers the
nine C-ABI calls with the CPU oracle instead of the CUDA engine (TEST INFRASTRUCTURE; libmpmc_b200.so has no CPU path).  What
is under test here is everything between `energy(system_t*)` and the C ABI: the list walks and the one-pass cached-sequence
check, change detection (set_positions ranges vs. set_atoms), uVT insert / remove / restore relinking, the observables
epilogue (energy.c:146-223) and the write-back -- against the reference's own fixed-seed trajectories."""
import os
import subprocess

import pytest

from tests import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STUB_DIR = os.path.join(ROOT, "tests", "hoststub")
PROGRAM = os.path.join(STUB_DIR, "_build", "mpmc_hoststub")
SHIM_OBJ = os.path.join(ROOT, "mpmc_b200", "shim", "_build", "obj", "energy_shim.o")


@pytest.fixture(scope="module")
def hoststub():
    if not os.path.exists(SHIM_OBJ):
        pytest.skip("the shim objects are built where /root/reference is present (__graft_entry__.build())")
    r = subprocess.run(["make", "-s", "-C", STUB_DIR], capture_output=True, text=True)
    assert r.returncode == 0 and os.path.exists(PROGRAM), r.stdout[-2000:] + r.stderr[-2000:]
    return PROGRAM


@pytest.mark.parametrize("host_path", ["cached_sequence", "list_walks", "threaded"])
@pytest.mark.parametrize("name", ["traj_small_nvt", "traj_small_uvt", "traj_co2_nvt", "traj_co2_uvt", "traj_co2_npt",
                                  "traj_co2_uvt_cavity", "traj_co2_uvt_insert_input"])
def test_shim_host_logic_reproduces_reference_trajectory(hoststub, name, host_path, tmp_path):
    """traj_co2_npt: volume moves -- every accepted or reverted change of the cell must reach set_params (its GPU run:
    tests/test_gpu_shim.py).  `threaded`: the one pass and the scatter into atom_t with 4 host threads (the default above
    4,096 atoms, forced here on the small systems), on a group of two devices as the shim parses MPMC_B200_DEVICES."""
    # MPMC_B200_SHIM_CHECK=1: after every one-pass call the shim compares its cached sequence, atom_array and molecule_array
    # with a real walk of the lists and dies on any difference
    env = {"MPMC_B200_SHIM_SLOW": "1"} if host_path == "list_walks" else {"MPMC_B200_SHIM_CHECK": "1"}
    if host_path == "threaded":
        env.update({"MPMC_B200_HOST_THREADS": "4", "MPMC_B200_HOST_THREADS_MIN": "0", "MPMC_B200_DEVICES": "0-1"})
    else:
        env["MPMC_B200_HOST_THREADS"] = "1"
    gu.run_and_check_trajectory(hoststub, name, tmp_path, env)


@pytest.mark.parametrize("name,batch", [("traj_small_nvt", 8), ("traj_small_nvt", 64), ("traj_co2_nvt", 16), ("traj_co2_uvt", 8),
                                        ("traj_co2_npt", 8)])
def test_resident_steps_reproduce_reference_trajectory(hoststub, name, batch, tmp_path):
    """SURVEY 8f f3, MPMC_B200_RESIDENT_STEPS: the shim reads the driver's next random numbers ahead, has a batch of NVT
    displacement steps moved, evaluated and decided behind the C ABI (mpmc_b200_mc_steps; here the CPU stand-in, whose move
    arithmetic is the plain-C twin of mc_step_kernel.cu) and answers the following energy() calls from the batch after
    checking the trial coordinates make_move() really produced, bit for bit.  The unmodified mc() must see exactly the
    reference's trajectory -- every energy row, the final configuration -- and every NVT step after the first of a batch
    must have been answered from the batch (none dropped); in NPT the batches end at the volume moves; uVT runs are not
    eligible and stay on the normal path."""
    env = {"MPMC_B200_RESIDENT_STEPS": str(batch), "MPMC_B200_SHIM_CHECK": "1", "MPMC_B200_SHIM_TIMING": "1",
           "MPMC_B200_HOST_THREADS": "1"}
    r = gu.run_and_check_trajectory(hoststub, name, tmp_path, env)
    import json
    import re

    steps = int(re.findall(r"^numsteps\s+(\d+)", json.load(open(os.path.join(gu.GOLD, name + ".json")))["input"], re.M)[-1])
    m = re.search(r"MPMC_B200_RESIDENT_STEPS: (\d+) batches, (\d+) steps answered from a batch, (\d+) batches cut short", r.stderr)
    if "nvt" in name:
        assert m, r.stderr[-2000:]
        batches, served, dropped = map(int, m.groups())
        assert served == steps and dropped == 0 and batches == -(-steps // batch), (batches, served, dropped, steps)
    elif "npt" in name:  # the displacement steps between volume moves are served; a batch ends where a volume move comes
        assert m, r.stderr[-2000:]
        batches, served, dropped = map(int, m.groups())
        assert served > steps // 2 and dropped == 0, (batches, served, dropped, steps)
    else:
        assert m is None or int(m.group(2)) == 0


DRIVER = os.path.join(STUB_DIR, "_build", "shim_driver_hoststub")


def _single_point_cases():
    from tests.test_gpu_shim import SINGLE_POINT  # the same list the GPU drop-in test walks

    return SINGLE_POINT


@pytest.mark.parametrize("name", _single_point_cases())
def test_shim_option_plumbing_single_point(hoststub, name, tmp_path):
    """energy(system_t*) through the shim for every option family (solver switches, Wolf / Ewald fields, full-Ewald
    polarization, mixing rules, FH, autoreject, rd_only, Wolf electrostatics, rd_crystal, the PCN-61 framework): what
    shim_fill_params() hands over and what the epilogue and the write-back leave in system_t / atom_t must equal the
    fixture generated from the unmodified reference -- here with the oracle behind the C ABI, on the GPU in
    tests/test_gpu_shim.py::test_dropin_energy_matches_reference_fixture."""
    import numpy as np

    from mpmc_b200 import systems
    from oracle import oracle  # snapshot parser

    s, exp = gu.load(name)
    first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
    s.meta["atom_mass"] = np.where(first, s.mol_mass, 0.0)
    inp = systems.write_reference_case(s, str(tmp_path))
    r = subprocess.run([DRIVER, os.path.basename(inp)], cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "MPMC_SNAPSHOT" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    _, got, _ = oracle.parse_snapshot(r.stdout, name)
    assert got["returned"] == got["total"] or got["total"] >= 1e40 or exp.get("total", 0) >= 1e40
    for k in ("rd", "coulombic", "polar", "total"):
        b = exp[k]
        tol = 1e-10 * abs(b)
        if k in ("coulombic", "total"):
            tol += 1e-13 * max(abs(exp.get("es_real", 0.0)), abs(exp.get("es_self", 0.0))) + 1e-10 * (abs(exp["rd"]) + abs(exp["polar"])) * (k == "total")
        assert abs(got[k] - b) <= tol, f"{name} {k}: shim + oracle {got[k]!r} vs reference {b!r}"
    assert int(got["polar_iterations"]) == int(exp["polar_iterations"])
    assert int(got["iter_success"]) == int(exp["iter_success"])
    if s.params.polarization and not s.params.rd_only and exp["total"] < 1e40:
        gu.assert_atom_arrays(got, exp, s, rtol=1e-10, label=name)


@pytest.mark.parametrize("name,calls_expected,fast_min", [("traj_co2_nvt", 301, 300), ("traj_co2_uvt", 301, 290),
                                                         ("traj_small_uvt", 301, 290)])
def test_cached_sequence_survives_relinking(hoststub, name, calls_expected, fast_min, tmp_path):
    """The MC driver relinks the lists all the time: a rejected displacement links the backup copy where the altered molecule
    was (restore(), mc_moves.c:789-797), uVT inserts and removes molecules and undoes that on rejection.  The one-pass host
    path edits its cached pointer sequence in place (shim_resync: removal / insertion / replacement of one node at a time)
    instead of falling back to the full list walks; MPMC_B200_SHIM_CHECK=1 compares the result with a real walk every call."""
    import json
    import re

    d = json.load(open(os.path.join(gu.GOLD, name + ".json")))
    (tmp_path / d["pqr_name"]).write_text(d["pqr"])
    (tmp_path / "input").write_text(d["input"])
    r = subprocess.run([hoststub, "input"], cwd=str(tmp_path), capture_output=True, text=True, timeout=900,
                       env=dict(os.environ, MPMC_B200_SHIM_TIMING="1", MPMC_B200_SHIM_CHECK="1"))
    assert r.returncode == 0, r.stderr[-2000:]
    m = re.search(r"(\d+) energy\(\) calls \((\d+) on the cached-sequence pass\)", r.stderr)
    assert m, r.stderr[-2000:]
    calls, fast = int(m.group(1)), int(m.group(2))
    assert calls == calls_expected and fast >= fast_min, f"{name}: {fast} of {calls} calls on the cached-sequence pass"


@pytest.mark.parametrize("extra,needle", [
    ("cdvdw_9th_repulsion on\n", "cdvdw repulsion"),
    ("ensemble surf\nsurf_min 2.0\nsurf_max 4.0\nsurf_inc 0.5\nsurf_ang 0.7\n", "surf"),
])
def test_shim_refuses_options_the_engine_does_not_implement(hoststub, extra, needle, tmp_path):
    """No silent fallback through the drop-in: cdvdw_9th_repulsion changes the LJ mixing inside pair_exclusions()
    (pairs.c:115-126) and the surf ensembles evaluate lj_nopbc() / coulombic_nopbc() over pair lists this build never
    allocates -- both must end in error() + die(), not in a run with wrong energies."""
    import json

    d = json.load(open(os.path.join(gu.GOLD, "traj_co2_nvt.json")))
    (tmp_path / d["pqr_name"]).write_text(d["pqr"])
    text = d["input"]
    if "ensemble" in extra:
        text = text.replace("ensemble nvt\n", "")
    (tmp_path / "input").write_text(text + "\n" + extra)
    r = subprocess.run([hoststub, "input"], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    out = r.stdout + r.stderr
    assert r.returncode != 0, out[-2000:]
    assert "not implemented by the B200 energy engine" in out and needle in out, out[-2000:]


def test_energy_no_observables_through_the_shim(hoststub, tmp_path):
    """energy_no_observables() (energy.c:232-318): the same energy, system->observables left alone."""
    import numpy as np

    from mpmc_b200 import systems
    from oracle import oracle

    s, exp = gu.load("mixed_default")
    first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
    s.meta["atom_mass"] = np.where(first, s.mol_mass, 0.0)
    inp = systems.write_reference_case(s, str(tmp_path))
    r = subprocess.run([DRIVER, os.path.basename(inp), "1", "0", "n"], cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "MPMC_SNAPSHOT" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    _, got, _ = oracle.parse_snapshot(r.stdout, "mixed_default")
    assert abs(got["returned"] - exp["total"]) <= 1e-10 * (abs(exp["rd"]) + abs(exp["polar"]) + abs(exp["coulombic"]))
    assert got["total"] == 0.0 and got["rd"] == 0.0 and got["polar"] == 0.0  # observables_t untouched (zero-allocated)
    gu.assert_atom_arrays(got, exp, s, rtol=1e-10, label="no_observables")
