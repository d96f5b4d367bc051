"""Pins of the benchmarked configurations (tests/golden/pin_*.npz): tied to the generators, and the smallest one
re-derived with the oracle on the CPU (the large ones take minutes: oracle/make_pins.py)."""
import numpy as np
import pytest

import bench
from tests import pins


@pytest.mark.parametrize("name", pins.names())
def test_pin_is_tied_to_the_generator(name):
    pin = pins.load(name)
    s = bench.build_workload(pin["meta"]["workload"])
    assert s.n == pin["meta"]["n"]
    assert pins.check_inputs(pin, s) == []
    assert pin["meta"]["pinned_by"] in ("oracle", "reference")
    for k in pins.ARRAYS:
        assert pin["sub_" + k].shape[0] == (s.n + pin["meta"]["stride"] - 1) // pin["meta"]["stride"]


def test_oracle_reproduces_the_sweep_pin():
    from oracle import oracle

    pin = pins.load("pin_h2_fluid_16000")
    s = bench.build_workload("h2_fluid_16000")
    e, _, arr = oracle.energy(s, nthreads=0)
    dev = pins.deviation(pin, e.to_dict(), arr)
    assert dev["iterations_equal"] and dev["max"] <= 1e-12, dev


def test_deviation_notices_a_wrong_answer():
    pin = pins.load("pin_h2_fluid_16000")
    e = dict(pin["energies"])
    assert pins.deviation(pin, e, None)["max"] == 0.0
    e["polar"] *= 1.0 + 1e-9
    assert pins.deviation(pin, e, None)["energies"] > 5e-10
    n = pin["meta"]["n"]
    arr = {k: np.zeros((n, 3)) for k in ("mu", "ef_static", "ef_induced", "ef_induced_change")}
    assert pins.deviation(pin, pin["energies"], arr)["arrays"] > 0.5
    # a permutation of rows that leaves the subsample rows alone is caught by the weighted checksum
    full = {k: np.zeros((n, 3)) for k in ("mu",)}
    full["mu"][::16] = pin["sub_mu"]
    d0 = pins.deviation(pin, pin["energies"], full)
    assert d0["arrays"] == 0.0 and d0["checksums"] > 1e-3


def test_reference_made_pin_and_the_reference_running_sum():
    """pin_co2_fluid_12k (BASELINE config 2, made by the UNMODIFIED reference): lj() adds every pair's (rd_energy + lrc)
    to ONE running sum (lj.c:256), which at 12,288 atoms carries ~2e-10 of rounding of its own.  The oracle in the
    reference's order reproduces the reference's rd BIT FOR BIT; its separated row-wise sums are the well-conditioned
    value the CUDA path is held to at 1e-10 (tests/test_gpu_pins.py)."""
    from oracle import oracle

    pin = pins.load("pin_co2_fluid_12k")
    assert pin["meta"]["pinned_by"] == "reference" and pin["meta"]["rd_reference_order_reproduced"]
    s = bench.build_workload("co2_fluid_12k")
    e1, _, _ = oracle.energy(s, nthreads=1)
    assert e1.rd == pin["energies"]["rd"]
    for k in ("es_real", "es_recip", "es_self"):
        assert abs(getattr(e1, k) - pin["energies"][k]) <= 1e-12 * abs(pin["energies"][k]), k
    e0, _, _ = oracle.energy(s, nthreads=0)
    assert e0.rd == pin["meta"]["rd_conditioned"]
    noise = abs(e0.rd - e1.rd) / abs(e1.rd)
    assert 1e-11 < noise < 1e-9  # the reference's own rounding at this size
    dev = pins.deviation(pin, e0.to_dict(), None)
    assert dev["energies"] <= 1e-10 and dev["rd_vs_reference"] <= 1.01 * dev["reference_rd_noise"] + 1e-12
