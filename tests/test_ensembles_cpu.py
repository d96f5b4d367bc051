"""The ensemble launcher (mpmc_b200/ensembles.py: independent fixed-seed chains, one per GPU, the reference's own
parallel mode -- main.c:50-52, mc.c:431-432) on the CPU: the chains here are the drop-in's own objects over the host
stand-in of the C ABI (tests/hoststub, TEST INFRASTRUCTURE), so what is under test is the launcher: seeding, one
directory and one device per chain, parsing of the MC driver's `OUTPUT:` averages (output.c:1133-1280), combination."""
import json
import os
import subprocess

import pytest

from mpmc_b200 import ensembles
from tests import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STUB_DIR = os.path.join(ROOT, "tests", "hoststub")
PROGRAM = os.path.join(STUB_DIR, "_build", "mpmc_hoststub")


@pytest.fixture(scope="module")
def hoststub():
    if not os.path.exists(os.path.join(ROOT, "mpmc_b200", "shim", "_build", "obj", "energy_shim.o")):
        pytest.skip("the shim objects are built where /root/reference is present (__graft_entry__.build())")
    r = subprocess.run(["make", "-s", "-C", STUB_DIR], capture_output=True, text=True)
    assert r.returncode == 0 and os.path.exists(PROGRAM), r.stdout[-2000:] + r.stderr[-2000:]
    return PROGRAM


def test_parse_averages():
    text = """OUTPUT: potential energy = -1234.50000 +- 12.25000 K
OUTPUT: N = 155.50000 +- 0.50000 molecules
OUTPUT: polarization energy = -10.12345 +- 0.00100 K (iterations = 10.0 +- 0.0 rrms = 0.0 +- 0.0)
OUTPUT: wt % = 1.23400 +- 0.00200 %
OUTPUT: qst = 6.12345 kJ/mol
OUTPUT: Completed step 10/10  (100.000 %)
"""
    a = ensembles.parse_averages(text)
    assert a["potential energy"] == {"value": -1234.5, "error": 12.25, "unit": "K"}
    assert a["N"]["value"] == 155.5 and a["wt %"]["value"] == 1.234 and a["qst"]["value"] == 6.12345
    assert a["polarization energy"]["value"] == -10.12345
    c = ensembles.combine([a, {k: dict(v, value=v["value"] + 2.0) for k, v in a.items()}])
    assert abs(c["N"]["value"] - 156.5) < 1e-12 and abs(c["N"]["error"] - 1.0) < 1e-12 and c["N"]["chains"] == 2


def test_seeded_input_and_devices():
    t = ensembles.seeded_input("ensemble nvt\npreset_seeds 5\nseed 9\nnumsteps 3\n", 42)
    assert t.count("preset_seeds") == 1 and "preset_seeds 42" in t and "seed 9" not in t and "numsteps 3" in t
    assert ensembles.parse_devices("0-2,5", 4) == [0, 1, 2, 5] and ensembles.parse_devices(None, 3) == [0, 1, 2]


def test_launcher_runs_independent_seeded_chains(hoststub, tmp_path):
    d = json.load(open(os.path.join(gu.GOLD, "traj_small_uvt.json")))
    src = tmp_path / "case"
    src.mkdir()
    (src / d["pqr_name"]).write_text(d["pqr"])
    (src / "input").write_text(d["input"].replace("numsteps 300", "numsteps 40"))
    r = ensembles.launch(str(src / "input"), 3, devices=[0, 1, 2], seeds=[777, 777, 12], workdir=str(tmp_path / "run"),
                         program=hoststub)
    assert [c["rc"] for c in r["chains"]] == [0, 0, 0], r
    assert [c["device"] for c in r["chains"]] == [0, 1, 2]
    a0, a1, a2 = (c["averages"] for c in r["chains"])
    assert "potential energy" in a0 and "N" in a0
    assert a0 == a1                      # same seed, same chain
    assert a0["potential energy"] != a2["potential energy"]  # another seed, another chain
    assert r["combined"]["N"]["chains"] == 3
    for k in range(3):
        assert os.path.exists(os.path.join(str(tmp_path / "run"), f"chain_{k}", "small.energy.dat"))
    # chain 0 (the golden seed) walks the reference's own trajectory
    rows = [l for l in open(os.path.join(str(tmp_path / "run"), "chain_0", "small.energy.dat")).read().splitlines() if not l.startswith("#")]
    for a, b in zip(rows, d["energy_rows"]):
        assert a.split()[0] == b.split()[0] and abs(float(a.split()[1]) - float(b.split()[1])) <= 2e-6 + 1e-9 * abs(float(b.split()[1]))
