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


@pytest.mark.parametrize("host_path", ["cached_sequence", "list_walks"])
@pytest.mark.parametrize("name", ["traj_small_nvt", "traj_small_uvt", "traj_co2_nvt", "traj_co2_uvt"])
def test_shim_host_logic_reproduces_reference_trajectory(hoststub, name, host_path, tmp_path):
    env = {"MPMC_B200_SHIM_SLOW": "1"} if host_path == "list_walks" else {}
    gu.run_and_check_trajectory(hoststub, name, tmp_path, env)
