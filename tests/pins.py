"""TEST / MEASUREMENT INFRASTRUCTURE (not part of the product package).

Pins of the benchmarked configurations (tests/golden/pin_<workload>.npz, written by oracle/make_pins.py).

A pin is the answer of the reference (where it can load the system) or of the reference-pinned oracle for one of
bench.py's workloads at its FULL size: energies by component, iteration count, a 1-in-16 subsample of every per-atom
array polar() leaves in atom_t, three whole-array checksums per array, and SHA-256 digests of the inputs that tie the
pin to the generator in mpmc_b200.systems.  Data only: nothing here imports or runs the oracle, so bench.py can report
`config.parity` on the GPU box and the -m gpu tests can hold the benchmarked runs to 1e-10.
"""
from __future__ import annotations

import hashlib
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ARRAYS = ("mu", "ef_static", "ef_induced", "ef_induced_change", "dipole_rrms", "rank_metric")
INPUTS = ("pos", "charge", "polarizability", "epsilon", "sigma", "mol_mass", "mol_id", "frozen")
ENERGY_KEYS = ("rd", "es_real", "es_recip", "es_self", "coulombic", "polar", "total")
STRIDE = 16


def row_weights(n: int) -> np.ndarray:
    """Fixed weights in [0.5, 1.5): a Weyl sequence of the row index (exact integer arithmetic, reproducible)."""
    i = np.arange(n, dtype=np.uint64)
    return 0.5 + ((i * np.uint64(2654435761)) % np.uint64(1 << 20)).astype(np.float64) / float(1 << 20)


def checksums(a: np.ndarray) -> list:
    """[sum, sum of squares, row-weighted sum, sum of magnitudes] of a per-atom array ([n] or [n][3])."""
    a = np.asarray(a, dtype=np.float64)
    w = row_weights(a.shape[0])
    if a.ndim == 2:
        w = w[:, None]
    return [float(np.sum(a)), float(np.sum(a * a)), float(np.sum(w * a)), float(np.sum(np.abs(a)))]


def input_digests(s) -> dict:
    d = {k: hashlib.sha256(np.ascontiguousarray(getattr(s, k)).tobytes()).hexdigest() for k in INPUTS}
    d["params"] = hashlib.sha256(json.dumps(s.params.to_dict(), sort_keys=True).encode()).hexdigest()
    return d


def path(name: str) -> str:
    return os.path.join(GOLD, name + ".npz")


def names() -> list:
    import glob

    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLD, "pin_*.npz")))


def for_workload(workload: str):
    """Name of the pin of a bench.py workload, or None."""
    p = "pin_" + workload
    return p if os.path.exists(path(p)) else None


def load(name: str) -> dict:
    z = np.load(path(name))
    pin = {"name": name, "energies": json.loads(str(z["expected_json"])), "meta": json.loads(str(z["meta_json"]))}
    for k in ARRAYS:
        pin["sub_" + k] = z["sub_" + k]
        pin["sum_" + k] = z["sum_" + k]
    return pin


def check_inputs(pin: dict, s) -> list:
    """Names of the inputs whose digest differs from the pinned system (empty = the same system)."""
    want = pin["meta"]["digests"]
    got = input_digests(s)
    return [k for k in want if got.get(k) != want[k]]


def deviation(pin: dict, energies: dict, arrays: dict | None) -> dict:
    """Largest deviations of a run from the pin.  Energies: relative to the pinned value, the Ewald total and the
    grand total against the scale of their parts (they are differences of much larger pieces; see
    tests/golden_util.py).  Arrays: subsample rows against the array's largest magnitude; checksums against the
    pinned sum of magnitudes (sum, weighted sum) or sum of squares."""
    exp = pin["energies"]
    out = {"energies": 0.0, "arrays": 0.0, "checksums": 0.0, "iterations_equal": True, "worst": ""}

    def note(kind, v, what):
        if v > out[kind]:
            out[kind] = float(v)
        if v >= max(out["energies"], out["arrays"], out["checksums"]):
            out["worst"] = what

    es_scale = max(abs(exp.get("es_real", 0.0)), abs(exp.get("es_self", 0.0)), abs(exp.get("es_recip", 0.0)))
    # Reference-made pins of 10^4-atom systems: lj() adds every pair's (rd_energy + lrc) to ONE running sum (lj.c:256),
    # which at this size carries ~1e-10 of rounding of its own.  make_pins.py proves that (the oracle in the
    # reference's order reproduces the reference's rd bit for bit) and stores the well-conditioned sum beside it:
    # rd and total are held to 1e-10 against THAT, and to 1e-10 + the reference's measured noise against the reference.
    rd_cond = pin["meta"].get("rd_conditioned")
    if rd_cond is not None and "rd" in energies and exp.get("rd"):
        out["reference_rd_noise"] = abs(exp["rd"] - rd_cond) / abs(exp["rd"])
        out["rd_vs_reference"] = abs(energies["rd"] - exp["rd"]) / abs(exp["rd"])
    for k in ENERGY_KEYS:
        if k not in exp or k not in energies:
            continue
        b = exp[k]
        if rd_cond is not None and k == "rd":
            b = rd_cond
        if rd_cond is not None and k == "total":
            b = (exp["total"] - exp["rd"]) + rd_cond
        scale = abs(b)
        if k == "coulombic":
            scale = max(scale, 1e-3 * es_scale)
        if k == "total":
            scale = max(scale, abs(exp.get("rd", 0.0)) + abs(exp.get("polar", 0.0)), 1e-3 * es_scale)
        if scale > 0.0:
            note("energies", abs(energies[k] - b) / scale, k)
    if "polar_iterations" in exp and "polar_iterations" in energies:
        out["iterations_equal"] = int(exp["polar_iterations"]) == int(energies["polar_iterations"])
    if arrays is not None:
        stride = int(pin["meta"].get("stride", STRIDE))
        for k in ARRAYS:
            if k not in arrays:
                continue
            a = np.asarray(arrays[k], dtype=np.float64)
            sub, sums = pin["sub_" + k], pin["sum_" + k]
            scale = float(np.abs(sub).max())
            if k == "ef_induced_change":  # a difference of two O(|ef_induced|) sums
                scale = max(scale, float(np.abs(pin["sub_ef_induced"]).max()))
            if k == "dipole_rrms":  # a ratio of small numbers: reported on its own, held to 1e-8 by the tests
                out["rrms"] = float(np.abs(a[::stride] - sub).max()) / max(1.0, scale)
                continue
            if scale > 0.0:
                note("arrays", float(np.abs(a[::stride] - sub).max()) / scale, k)
            elif a.size:
                note("arrays", float(np.abs(a).max()), k)
            got = checksums(a)
            mag = float(sums[3])
            if mag > 0.0:
                note("checksums", abs(got[0] - sums[0]) / mag, k + ".sum")
                note("checksums", abs(got[2] - sums[2]) / mag, k + ".wsum")
            if sums[1] > 0.0:
                note("checksums", abs(got[1] - sums[1]) / float(sums[1]), k + ".sum2")
    out["max"] = max(out["energies"], out["arrays"], out["checksums"])
    return out
