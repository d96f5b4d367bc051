#!/usr/bin/env python
"""profiles/mc_dropin_bench.py -- MC steps/s through the DROP-IN BINARY (BASELINE.json's first metric).

The reference's unmodified host program (parser, MC driver, Mersenne RNG, writers) linked against
mpmc_b200/shim/energy_shim.c + libmpmc_b200.so (mpmc_b200/shim/_build/mpmc_b200) runs fixed-seed NVT / uVT
trajectories of synthetic systems shaped like BASELINE's configs 2, 3 and 5; steps/s = (steps_long - steps_short) /
(wall_long - wall_short), so parsing, the initial energy() and process start cancel.  Where the unmodified reference
(oracle/_ref/mpmc_ref, test infrastructure / CPU baseline only) can hold the system in memory it runs the same
input for a few steps on one host core.

    python profiles/mc_dropin_bench.py [--quick] > gpurun_out/mc_dropin.jsonl
"""
from __future__ import annotations

import json
import os

import numpy as np
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mpmc_b200 import systems  # noqa: E402

DROPIN = os.path.join(ROOT, "mpmc_b200", "shim", "_build", "mpmc_b200")
REFERENCE = os.path.join(ROOT, "oracle", "_ref", "mpmc_ref")


def run(binary, s, ensemble, steps, extra, timeout, env=None):
    with tempfile.TemporaryDirectory() as td:
        ex = f"numsteps {steps}\ncorrtime {max(steps, 1)}\npreset_seeds 12345\n" + extra
        systems.write_reference_case(s, td, ensemble=ensemble, extra=ex)
        t0 = time.perf_counter()
        try:
            r = subprocess.run([binary, "input"], cwd=td, capture_output=True, text=True, timeout=timeout,
                               env=dict(os.environ, **(env or {})))
        except subprocess.TimeoutExpired:
            return None, "timeout"
        dt = time.perf_counter() - t0
        if r.returncode != 0:
            return None, (r.stdout[-300:] + r.stderr[-300:]).replace("\n", " | ")
        return dt, ""


def steps_per_s(binary, s, ensemble, short, long_, extra, timeout, env=None):
    # process start + CUDA context creation (~1 s, +-0.3 s) cancel in the difference; the long run is sized so that
    # the difference is several seconds
    a, ea = run(binary, s, ensemble, short, extra, timeout, env)
    if a is None:
        return None, ea
    a2, _ = run(binary, s, ensemble, short, extra, timeout, env)
    if a2 is not None:
        a = min(a, a2)
    b, eb = run(binary, s, ensemble, long_, extra, timeout, env)
    if b is None:
        return None, eb
    return (long_ - short) / max(b - a, 1e-9), ""


def main():
    quick = "--quick" in sys.argv
    uvt = "pressure 70.0\nh2_fugacity on\ninsert_probability 0.666\nmove_factor 0.001\nrot_factor 0.01\n"
    nvt = "move_factor 0.01\nrot_factor 0.01\n"
    from tests import golden_util as gu

    socmof, _ = gu.load("ref_socmof")  # the reference's shipped In-soc-MOF + 156 H2 system (1,236 atoms) as a flat fixture
    first = np.r_[True, socmof.mol_id[1:] != socmof.mol_id[:-1]]
    socmof.meta["atom_mass"] = np.where(first, socmof.mol_mass, 0.0)  # molecule masses on the first atom (as test_gpu_shim does)
    cases = [
        # name, system, ensemble, extra, (short, long) steps for the drop-in, (short, long) for the reference or None
        ("socMOF+BSSP 1,236 atoms NVT, FH4, Jacobi x10 -- the system of the reference's only published timing "
         "(0.478-0.492 s/step on a Xeon E5-2670 core, 0.136 s/step with its fp32 CUDA kernel; BASELINE.md 1)", socmof, "nvt", nvt,
         (100, 4100), (2, 12)),
        ("h2_fluid_1000 NVT, Jacobi x10, FH4 (config 5, small N)", systems.h2_fluid(1000), "nvt", nvt, (100, 6100), (2, 42)),
        ("co2_fluid_12k NVT, LJ+Ewald, no polarization (config 2)", systems.co2_fluid(4096), "nvt", nvt, (50, 1550), (1, 3)),
        ("mof_h2_21k uVT, Jacobi x4, FH4, cutoff 8 A (config 3 shape: 128.4 x 42.8 x 42.8 A)",
         systems.mof_h2(cells=(9, 3, 3), polar_max_iter=4), "uvt", uvt, (20, 820), None),
        ("mof_h2_63k NVT, Jacobi x10 (config 4 system as a trajectory)", systems.mof_h2(), "nvt", nvt, (5, 125), None),
    ]
    if "--only-tolerance" in sys.argv:  # SURVEY 8d M4 in tolerance mode (the reference's production setting) as a trajectory
        tol = systems.mof_h2()
        tol.params.polar_max_iter, tol.params.polar_precision = 0, 1e-5
        cases = [("mof_h2_63k NVT, Jacobi to polar_precision 1e-5 (config 4, tolerance mode)", tol, "nvt", nvt, (5, 65), None)]
    if quick:
        cases = cases[:1]
    if "--only-socmof" in sys.argv:
        cases = cases[:1]
    if "--only-co2" in sys.argv:  # config 2 alone (with --incremental: the opt-in incremental path beside the default)
        cases = cases[2:3]
    for name, s, ens, extra, (a, b), ref in cases:
        v, err = steps_per_s(DROPIN, s, ens, a, b, extra, 900)
        line = {"case": name, "n_atoms": int(s.n), "ensemble": ens, "dropin_mc_steps_per_s": v, "dropin_error": err or None,
                "dropin_steps": [a, b]}
        if "--ab" in sys.argv:  # the reference-shaped list walks instead of the cached-sequence pass
            v2, err2 = steps_per_s(DROPIN, s, ens, a, b, extra, 900, {"MPMC_B200_SHIM_SLOW": "1"})
            line["dropin_mc_steps_per_s_reference_shaped_walks"] = v2
        if "--incremental" in sys.argv:  # only the moved molecule's pairs and S(k) terms per step
            v3, err3 = steps_per_s(DROPIN, s, ens, a, b, extra, 900, {"MPMC_B200_INCREMENTAL": "1"})
            line["dropin_mc_steps_per_s_incremental"] = v3
            line["incremental_error"] = err3 or None
        if "--warm-start" in sys.argv:  # tolerance-mode solve started from the previous step's dipoles
            v4, err4 = steps_per_s(DROPIN, s, ens, a, b, extra, 900, {"MPMC_B200_WARM_START": "1"})
            line["dropin_mc_steps_per_s_warm_start"] = v4
            line["warm_start_error"] = err4 or None
        if ref is not None and os.path.exists(REFERENCE) and "--no-reference" not in sys.argv:
            rv, rerr = steps_per_s(REFERENCE, s, ens, ref[0], ref[1], extra, 600)
            line.update({"reference_mc_steps_per_s_1core": rv, "reference_error": rerr or None, "reference_steps": list(ref)})
            if v and rv:
                line["ratio"] = v / rv
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
