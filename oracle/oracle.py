"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY (see oracle/mpmc_oracle.c header).

ctypes binding of the CPU restatement (oracle/liboracle.so) plus the parser for
the full-precision snapshots printed by oracle/ref_driver.c.  May be imported
only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Parity status: pinned (tests/test_oracle_golden.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from mpmc_b200.api import Counters, Energies, Params, System, MIX_C6, MIX_HALGREN, MIX_LB, MIX_WH

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "liboracle.so")
REF_DRIVER = os.path.join(_HERE, "_ref", "ref_driver")
REF_BINARY = os.path.join(_HERE, "_ref", "mpmc_ref")
_lib = None


def build(force: bool = False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(_HERE, "mpmc_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
    if os.path.exists("/root/reference/src/main/main.c"):
        subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        sig = [C.POINTER(Params), C.c_int, dp, dp, dp, dp, dp, dp, ip, ip, C.POINTER(Energies), C.POINTER(Counters),
               dp, dp, dp, dp, dp, dp, C.c_int]
        _lib.mpmc_oracle_energy.argtypes = sig
        _lib.mpmc_oracle_polar.argtypes = sig
        _lib.mpmc_oracle_energy.restype = C.c_int
        _lib.mpmc_oracle_polar.restype = C.c_int
        _lib.mpmc_oracle_max_threads.restype = C.c_int
    return _lib


def max_threads() -> int:
    return int(_load().mpmc_oracle_max_threads())


def _call(fn, s: System, nthreads: int):
    n = s.n
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
    out = {k: np.zeros((n, 3)) for k in ("mu", "ef_static", "ef_induced", "ef_induced_change")}
    out["dipole_rrms"] = np.zeros(n)
    out["rank_metric"] = np.zeros(n)
    e, c = Energies(), Counters()
    rc = fn(C.byref(s.params), n, dp(s.pos), dp(s.charge), dp(s.polarizability), dp(s.epsilon), dp(s.sigma),
            dp(s.mol_mass), ip(s.mol_id), ip(s.frozen), C.byref(e), C.byref(c), dp(out["mu"]), dp(out["ef_static"]),
            dp(out["ef_induced"]), dp(out["ef_induced_change"]), dp(out["dipole_rrms"]), dp(out["rank_metric"]),
            int(nthreads))
    if rc:
        raise RuntimeError(f"oracle: unsupported option combination (code {rc})")
    return e, c, out


def energy(s: System, nthreads: int = 0):
    """CPU restatement of energy(); nthreads=1 reproduces the reference's summation order."""
    return _call(_load().mpmc_oracle_energy, s, nthreads)


def polar(s: System, nthreads: int = 0):
    return _call(_load().mpmc_oracle_polar, s, nthreads)


# ---------------------------------------------------------------------------
# snapshots printed by ref_driver.c
# ---------------------------------------------------------------------------
def parse_snapshot(text: str, name: str = ""):
    """-> (System, expected dict with energies + per-atom arrays, timing seconds)."""
    lines = text.splitlines()
    start = max(i for i, l in enumerate(lines) if l.startswith("MPMC_SNAPSHOT"))
    par, en, atoms, timing = {}, {}, [], None
    for l in lines[start + 1:]:
        t = l.split()
        if not t:
            continue
        if t[0] == "param":
            vals = [float(x) for x in t[2:]]
            par[t[1]] = vals[0] if len(vals) == 1 else vals
        elif t[0] == "energy":
            en[t[1]] = float(t[2])
        elif t[0] == "timing":
            timing = float(t[2])
        elif t[0] == "atom":
            atoms.append([float(x) for x in t[1:]])
        elif t[0] == "END":
            break
    a = np.array(atoms, dtype=np.float64).reshape(len(atoms), -1)
    d = {}
    for k in ("cutoff", "volume", "ewald_alpha", "temperature", "cavity_autoreject_scale", "polar_damp",
              "polar_precision", "polar_gamma", "polar_wolf_alpha", "polar_wolf_alpha_lookup_cutoff",
              "polar_ewald_alpha"):
        d[k] = par[k]
    for k in ("ewald_kmax", "feynman_hibbs", "feynman_hibbs_order", "rd_lrc", "rd_only", "cavity_autoreject_absolute",
              "polarization", "polar_iterative", "damp_type", "polar_max_iter", "polar_gs", "polar_gs_ranked",
              "polar_sor", "polar_esor", "polar_palmo", "polar_rrms", "polar_zodid", "polar_ewald", "polar_ewald_full", "polar_wolf",
              "polar_wolf_full", "polar_wolf_alpha_lookup"):
        d[k] = int(par[k])
    d["basis"] = np.array(par["basis"]).reshape(3, 3)
    d["reciprocal_basis"] = np.array(par["reciprocal_basis"]).reshape(3, 3)
    d["es_wolf"] = int(par.get("wolf", 0))
    d["rd_crystal"] = int(par.get("rd_crystal", 0))
    d["rd_crystal_order"] = int(par.get("rd_crystal_order", 0))
    d["mixing"] = MIX_WH if par["waldmanhagler"] else MIX_HALGREN if par["halgren_mixing"] else MIX_C6 if par[
        "c6_mixing"] else MIX_LB
    p = Params.from_dict(d)
    sysm = System(p, a[:, 2:5], a[:, 5], a[:, 6], a[:, 7], a[:, 8], a[:, 10], a[:, 0].astype(np.int32),
                  a[:, 1].astype(np.int32), name)
    exp = dict(en)
    exp["polar_ewald_full"] = int(par.get("polar_ewald_full", 0))
    exp["ef_static"] = a[:, 11:14].copy()
    exp["ef_induced"] = a[:, 14:17].copy()
    exp["ef_induced_change"] = a[:, 17:20].copy()
    exp["mu"] = a[:, 20:23].copy()
    exp["dipole_rrms"] = a[:, 23].copy()
    exp["rank_metric"] = a[:, 24].copy()
    return sysm, exp, timing


def run_reference(input_path: str, repeat: int = 1, timeout: float = 3600.0):
    """Run the unmodified reference (oracle/_ref/ref_driver) on an MPMC input file."""
    if not os.path.exists(REF_DRIVER):
        raise FileNotFoundError(REF_DRIVER)
    wd = os.path.dirname(os.path.abspath(input_path))
    r = subprocess.run([REF_DRIVER, os.path.basename(input_path), str(repeat)], cwd=wd, capture_output=True, text=True,
                       timeout=timeout)
    if r.returncode != 0 or "MPMC_SNAPSHOT" not in r.stdout:
        raise RuntimeError(f"ref_driver failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
    return parse_snapshot(r.stdout, os.path.basename(wd))
