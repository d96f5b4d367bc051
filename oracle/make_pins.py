#!/usr/bin/env python
"""oracle/make_pins.py -- TEST INFRASTRUCTURE ONLY.

Pins for the configurations the benchmark numbers are quoted on (BASELINE.json configs 2, 4, 5 at their stated sizes):
tests/golden/pin_<workload>.npz.  The systems are too large to commit whole (7 MB of per-atom results at 64k atoms), so
a pin holds

  * the energies by component, the iteration count and the work counters;
  * a deterministic 1-in-`STRIDE` subsample of every per-atom array energy() / polar() leaves in atom_t
    (mu, ef_static, ef_induced, ef_induced_change, dipole_rrms, rank_metric), rows i = 0, STRIDE, 2 STRIDE, ...;
  * three checksums of each WHOLE array (sum, sum of squares, and a position-weighted sum whose weights are a fixed
    function of the row index, so that a permutation of rows does not go unnoticed);
  * SHA-256 of every input array, so that the pin is tied to the generator (mpmc_b200.systems): if the generator ever
    produced a different system the tests fail on the digest, not on a puzzling energy difference.

Who computed the expected values:
  pin_co2_fluid_12k          the UNMODIFIED reference (oracle/_ref/ref_driver; 12,288 atoms, ~15 GB of pair lists)
  pin_mof_h2_63k[...]        the oracle (oracle/mpmc_oracle.c, pinned to the reference by tests/test_oracle_golden.py):
  pin_h2_fluid_16000         the reference needs ~290 GB / ~45 GB for these systems

    python oracle/make_pins.py [name ...]      (minutes per case on 8 cores)
"""
from __future__ import annotations

import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mpmc_b200 import systems  # noqa: E402
from oracle import oracle  # noqa: E402

from tests.pins import ARRAYS, GOLD, INPUTS, STRIDE, checksums, input_digests  # noqa: E402


def build(name: str):
    """The same builder bench.py uses (imported lazily: bench.py lives at the repo root)."""
    import bench

    return bench.build_workload(name)


CASES = {
    # name: (workload of bench.py, who computes the expected values)
    "pin_mof_h2_63k": ("mof_h2_63k", "oracle"),
    "pin_co2_fluid_12k": ("co2_fluid_12k", "reference"),
    "pin_h2_fluid_16000": ("h2_fluid_16000", "oracle"),
    "pin_mof_h2_63k_gs_production": ("mof_h2_63k_gs_production", "oracle"),
    "pin_mof_h2_21k": ("mof_h2_21k", "oracle"),
}


def make(name: str):
    workload, who = CASES[name]
    s = build(workload)
    t0 = time.time()
    if who == "reference":
        first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
        keep = s.meta.get("atom_mass")
        with tempfile.TemporaryDirectory() as td:
            inp = systems.write_reference_case(s, td)
            parsed, exp, _ = oracle.run_reference(inp, timeout=6 * 3600.0)
        for k in INPUTS:
            assert np.array_equal(np.asarray(getattr(parsed, k)), np.asarray(getattr(s, k))), f"{name}: reference parsed {k} differently"
        del first, keep
        energies = {k: v for k, v in exp.items() if not isinstance(v, np.ndarray)}
        arr = {k: exp[k] for k in ARRAYS}
        counters = {}
        # lj() adds every pair's (rd_energy + lrc) to ONE running sum (lj.c:256); at this size that sum carries ~1e-10 of
        # rounding of its own.  The oracle in the reference's order (nthreads=1) must reproduce the reference's rd bit
        # for bit; its separated, row-wise sums (nthreads=0) are the well-conditioned value the CUDA path is held to.
        e1, _, _ = oracle.energy(s, nthreads=1)
        assert e1.rd == energies["rd"], f"{name}: the oracle in reference order gives rd {e1.rd!r}, the reference {energies['rd']!r}"
        e0, _, _ = oracle.energy(s, nthreads=0)
        extra_meta = {"rd_conditioned": e0.rd, "rd_reference_order_reproduced": True}
    else:
        e, c, arr = oracle.energy(s, nthreads=0)
        energies = e.to_dict()
        counters = {k: float(getattr(c, k)) for k in ("pairs_lj", "pairs_es", "pairs_field", "pairs_dipole", "contractions")}
        extra_meta = {}
    meta = {"workload": workload, "pinned_by": who, "stride": STRIDE, "n": int(s.n), "counters": counters,
            "digests": input_digests(s), "seconds": round(time.time() - t0, 1)}
    meta.update(extra_meta)
    out = {"expected_json": np.array(json.dumps(energies)), "meta_json": np.array(json.dumps(meta))}
    for k in ARRAYS:
        a = np.asarray(arr[k], dtype=np.float64)
        out["sub_" + k] = a[::STRIDE].copy()
        out["sum_" + k] = np.array(checksums(a))
    os.makedirs(GOLD, exist_ok=True)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print(f"{name}: n={s.n} by {who} in {time.time() - t0:.0f} s  total={energies['total']!r} polar={energies.get('polar')!r} "
          f"iters={energies.get('polar_iterations')}", flush=True)


def main():
    oracle.build()
    want = sys.argv[1:] or list(CASES)
    for name in want:
        make(name)


if __name__ == "__main__":
    main()
