"""The configurations the benchmark numbers are quoted on, at their FULL sizes, against their pins
(tests/golden/pin_*.npz, oracle/make_pins.py): BASELINE config 4 (mof_h2_63k: the default dispatch -- symmetric K4,
182 super tiles, far-cluster damping skip on -- and the Gauss-Seidel production options), config 2 (co2_fluid_12k,
pinned by the UNMODIFIED reference), a config-5 sweep point (h2_fluid_16000) and the config-3 shape (mof_h2_21k).
Energies by component <= 1e-10, every per-atom array on a 1-in-16 subsample and through three whole-array checksums,
identical iteration counts.  Called through the C ABI."""
import numpy as np
import pytest

import bench
from tests import pins

pytestmark = pytest.mark.gpu

PINS = pins.names()


def test_pins_exist():
    for need in ("pin_mof_h2_63k", "pin_co2_fluid_12k", "pin_h2_fluid_16000"):
        assert need in PINS, f"{need} missing: python oracle/make_pins.py {need}"


@pytest.mark.parametrize("name", PINS)
def test_cuda_matches_pin(engine, name):
    pin = pins.load(name)
    s = bench.build_workload(pin["meta"]["workload"])
    assert pins.check_inputs(pin, s) == [], "the generator no longer produces the pinned system"
    engine.set_system(s)
    got = engine.energy()
    polar = bool(s.params.polarization and not s.params.rd_only)
    arr = engine.atom_arrays() if polar else None
    dev = pins.deviation(pin, got.to_dict(), arr)
    assert dev["iterations_equal"], f"{name}: {got.polar_iterations} iterations vs {pin['energies']['polar_iterations']}"
    assert dev["energies"] <= 1e-10, f"{name}: energies deviate by {dev['energies']:.3e} ({dev['worst']})"
    assert dev["arrays"] <= 1e-10, f"{name}: per-atom arrays deviate by {dev['arrays']:.3e} ({dev['worst']})"
    assert dev["checksums"] <= 1e-10, f"{name}: array checksums deviate by {dev['checksums']:.3e} ({dev['worst']})"
    assert dev.get("rrms", 0.0) <= 1e-8
    if "rd_vs_reference" in dev:  # reference-made pin: see tests/pins.py on the rounding of lj()'s running sum
        assert dev["rd_vs_reference"] <= 1e-10 + 1.01 * dev["reference_rd_noise"], dev
    if polar and pin["meta"]["workload"] == "mof_h2_63k":
        # the dispatch the bench line is quoted on: symmetric contraction (>= 3,072 padded polarizable atoms)
        c = engine.counters()
        assert c.n_polarizable >= 8192 and engine.uses_symmetric_contraction()
    # a second evaluation of the resident system reproduces the first bit for bit (fixed-order sums, no atomics)
    again = engine.energy()
    assert again.total == got.total and again.polar == got.polar
    if polar:
        arr2 = engine.atom_arrays()
        assert np.array_equal(arr2["mu"], arr["mu"])


def test_contraction_in_rounds_under_a_small_partial_buffer_cap(tmp_path):
    """MPMC_B200_SYM_PART_MB bounds the J-side partial rows of the symmetric contraction (N^2/512 doubles for a whole
    contraction): above the cap a contraction is cut into launches over slice ranges that reuse one buffer.  Forced here
    with a 2 MB cap on the 21k-atom MOF (31 offsets -> several rounds) in a fresh process (the cap is read once), against
    the pin; on >= 2 GPUs also sharded over a device group."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r"""
import sys
sys.path.insert(0, %r)
import bench
from mpmc_b200 import Engine
from tests import pins
import torch
pin = pins.load("pin_mof_h2_21k")
s = bench.build_workload("mof_h2_21k")
devs = [None] + ([[0, 1]] if torch.cuda.device_count() >= 2 else [])
for d in devs:
    eng = Engine(0) if d is None else Engine(devices=d)
    eng.set_system(s)
    got = eng.energy()
    dev = pins.deviation(pin, got.to_dict(), eng.atom_arrays())
    assert dev["iterations_equal"] and dev["max"] <= 1e-10, (d, dev)
    again = eng.energy()
    assert again.polar == got.polar
    eng.close()
print("ROUNDS_OK", len(devs))
""" % root
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, MPMC_B200_SYM_PART_MB="2", MPMC_B200_SYM_MIN="0"))
    assert r.returncode == 0 and "ROUNDS_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
    # the other launch shape of the symmetric contraction: whole J super tiles per CTA (half tiles are the default since the
    # end of round 2, MPMC_B200_SYM_HSPLIT); both must meet the pin
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, MPMC_B200_SYM_HSPLIT="1", MPMC_B200_SYM_MIN="0"))
    assert r.returncode == 0 and "ROUNDS_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
    # the finest items (one j-cluster of a half tile per CTA; chosen automatically only for launches below ~3 waves)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, MPMC_B200_SYM_HSPLIT="8", MPMC_B200_SYM_MIN="0"))
    assert r.returncode == 0 and "ROUNDS_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
    # and the minimum image taken per pair everywhere (by default a cluster pair whose boxes prove one image for all its
    # atom pairs moves the i atoms once instead, MPMC_B200_SYM_UNISHIFT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, MPMC_B200_SYM_UNISHIFT="0", MPMC_B200_SYM_MIN="0"))
    assert r.returncode == 0 and "ROUNDS_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
