"""SURVEY 8f f4, writer half (mpmc_b200/shim/writers.c): the per-corrtime writers of the MC driver -- restart / final PQR
(write_molecules, src/io/output.c:315-564), trajectory (write_states, :674-852), molecular dipoles and fields (write_dipole /
write_field, :1029-1088) -- re-done as one list walk + per-thread formatting + a few fwrite calls.  Every file must be
BYTE FOR BYTE what the unmodified reference writes on the same input.  CPU test: the drop-in program with the oracle behind
the C ABI (tests/hoststub, test infrastructure) next to oracle/_ref/mpmc_ref, both run here."""
import os
import subprocess

import pytest

from mpmc_b200 import systems

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STUB_DIR = os.path.join(ROOT, "tests", "hoststub")
PROGRAM = os.path.join(STUB_DIR, "_build", "mpmc_hoststub")
SHIM_OBJ = os.path.join(ROOT, "mpmc_b200", "shim", "_build", "obj", "writers.o")
REFERENCE = os.path.join(ROOT, "oracle", "_ref", "mpmc_ref")
FILES = ["final.pqr", "restart.pqr", "restart.pqr.last", "traj.pdb", "dipole.dat", "field.dat"]


@pytest.fixture(scope="module")
def programs():
    if not os.path.exists(SHIM_OBJ) or not os.path.exists(REFERENCE):
        pytest.skip("needs the shim objects and oracle/_ref (built where /root/reference is present)")
    r = subprocess.run(["make", "-s", "-C", STUB_DIR], capture_output=True, text=True)
    assert r.returncode == 0 and os.path.exists(PROGRAM), r.stdout[-2000:] + r.stderr[-2000:]
    return PROGRAM, REFERENCE


def _case(which):
    if which == "h2_fluid_wrapped":
        s = systems.h2_fluid(600)
        extra = "wrapall on\n"
    elif which == "h2_fluid_long_output":  # %11.6f coordinates, un-wrapped positions, atom ids in the molecule column
        s = systems.h2_fluid(600)
        extra = "wrapall off\nlong_output on\nindependent_particle on\n"
    else:  # frozen framework + sorbate: F / M flags, molecules the dipole / field files skip
        s = systems.mof_h2(cells=(2, 2, 1))
        extra = "wrapall on\n"
    extra += ("numsteps 6\ncorrtime 2\nmove_factor 0.05\nrot_factor 0.5\npreset_seeds 20261105\ntraj_output traj.pdb\n"
              "dipole_output dipole.dat\nfield_output field.dat\n")
    return s, extra


@pytest.mark.parametrize("which", ["h2_fluid_wrapped", "h2_fluid_long_output", "mof_h2"])
def test_writers_are_byte_identical_to_the_reference(programs, which, tmp_path):
    program, reference = programs
    s, extra = _case(which)
    outs = {}
    for tag, exe, env in (("ref", reference, {}),
                          # lines formatted by 4 host threads (the threshold of 4,096 atoms lowered for the test)
                          ("buffered", program, {"MPMC_B200_HOST_THREADS": "4", "MPMC_B200_HOST_THREADS_MIN": "0",
                                                 "MPMC_B200_SHIM_TIMING": "1"}),
                          ("routed", program, {"MPMC_B200_HOST_THREADS": "1", "MPMC_B200_WRITERS": "reference",
                                               "MPMC_B200_SHIM_TIMING": "1"})):
        d = tmp_path / tag
        inp = systems.write_reference_case(s, str(d), ensemble="nvt", extra=extra)
        text = open(inp).read().replace("traj_output /dev/null\n", "").replace("wrapall on\n", "", 1)  # set by `extra`
        open(inp, "w").write(text)
        r = subprocess.run([exe, os.path.basename(inp)], cwd=str(d), capture_output=True, text=True, timeout=1500,
                           env=dict(os.environ, **env))
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        outs[tag] = r
    assert "(buffered, host threads)" in outs["buffered"].stderr and "write_states" in outs["buffered"].stderr
    assert "(reference)" in outs["routed"].stderr
    for f in FILES:
        want = (tmp_path / "ref" / f).read_bytes()
        assert len(want) > 0, f
        for tag in ("buffered", "routed"):
            got = (tmp_path / tag / f).read_bytes()
            assert got == want, f"{which}: {f} written by the drop-in ({tag}) differs from the reference's"
    # 3 corrtime blocks + the end-of-run write: the trajectory holds 3 frames of every atom
    assert (tmp_path / "buffered" / "traj.pdb").read_text().count("ENDMDL") == 3
