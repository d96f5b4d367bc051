"""Independent MC ensembles, one per GPU -- the reference's own parallel mode.

The reference runs `size` MPI ranks, each an independent Markov chain of the same input with its own seed, and the
root rank averages their observables (src/main/main.c:50-52, src/mc/mc.c:431-432, src/io/average.c:213-349).  Here the
same thing is N processes of the drop-in program `mpmc_b200 <input>` (the reference's unmodified MC driver over the CUDA
engine), chain k pinned to GPU k through MPMC_B200_DEVICE and seeded through `preset_seeds`; no collective is needed --
the chains never talk -- and the root's job is done here: parse every chain's final `OUTPUT:` averages and combine them
(mean over chains; standard error from the spread of the chain means, as average.c's sdom does across ranks).

    python -m mpmc_b200.ensembles <input file> --chains 8 [--devices 0-7] [--seeds 1,2,...] [--keep DIR]

Everything a chain writes (energy, trajectory, restart, dipole / field files) lands in its own directory chain_<k>/.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import time
from typing import Dict, List, Optional, Sequence

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "mpmc_b200", "shim", "_build", "mpmc_b200")

_NUM = r"([-+]?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?|-?nan|-?inf)"
_AVG = re.compile(r"^OUTPUT:\s+(.+?)\s+=\s+" + _NUM + r"\s+\+-\s+" + _NUM + r"\s*(\S.*)?$")
_ONE = re.compile(r"^OUTPUT:\s+(.+?)\s+=\s+" + _NUM + r"\s*(\S.*)?$")


def parse_averages(stdout: str) -> Dict[str, Dict[str, float]]:
    """The LAST averages block the MC driver printed (write_averages(), src/io/output.c:1133-1280):
    {"potential energy": {"value": v, "error": e, "unit": "K"}, "N": {...}, "wt %": {...}, ...}."""
    out: Dict[str, Dict[str, float]] = {}
    for line in stdout.splitlines():
        m = _AVG.match(line.strip())
        if m:
            out[m.group(1)] = {"value": float(m.group(2)), "error": float(m.group(3)), "unit": (m.group(4) or "").strip()}
            continue
        m = _ONE.match(line.strip())
        if m and m.group(1) in ("qst", "Energy/particle"):
            out[m.group(1)] = {"value": float(m.group(2)), "error": 0.0, "unit": (m.group(3) or "").strip()}
    return out


def parse_devices(spec: Optional[str], n: int) -> List[int]:
    if not spec:
        return list(range(n))
    dev: List[int] = []
    for part in spec.split(","):
        if "-" in part:
            a, b = part.split("-")
            dev.extend(range(int(a), int(b) + 1))
        elif part:
            dev.append(int(part))
    return dev


def host_threads_per_chain(chains: int) -> int:
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        cores = os.cpu_count() or 1
    return max(1, min(8, (cores - 1) // max(chains, 1)))


def seeded_input(text: str, seed: int) -> str:
    """The input with `preset_seeds <seed>` (src/io/input.c:118-122), replacing any seed line it had."""
    lines = [l for l in text.splitlines() if not re.match(r"^\s*(preset_seeds|seed)\b", l)]
    return "\n".join(lines) + f"\npreset_seeds {seed}\n"


def launch(input_path: str, chains: int, devices: Optional[Sequence[int]] = None, seeds: Optional[Sequence[int]] = None,
           workdir: Optional[str] = None, program: str = DROPIN, env: Optional[dict] = None, timeout: float = 24 * 3600.0):
    """Run `chains` independent chains of `input_path` (the files beside it are copied), chain k on devices[k % len].
    Returns {"chains": [per-chain dict], "combined": {...}, "wall_s": seconds}."""
    if not os.path.exists(program):
        raise FileNotFoundError(f"{program}: build the drop-in first (python -c 'import __graft_entry__ as g; g.build()')")
    src = os.path.dirname(os.path.abspath(input_path))
    text = open(input_path).read()
    devices = list(devices) if devices else list(range(chains))
    seeds = list(seeds) if seeds else list(range(1, chains + 1))
    own = workdir is None
    workdir = workdir or tempfile.mkdtemp(prefix="mpmc_ensembles_")
    procs = []
    t0 = time.perf_counter()
    for k in range(chains):
        d = os.path.join(workdir, f"chain_{k}")
        os.makedirs(d, exist_ok=True)
        for f in os.listdir(src):
            p = os.path.join(src, f)
            if os.path.isfile(p) and os.path.abspath(p) != os.path.abspath(input_path):
                shutil.copy(p, d)
        with open(os.path.join(d, "input"), "w") as f:
            f.write(seeded_input(text, seeds[k % len(seeds)]))
        e = dict(os.environ, **(env or {}))
        for v in ("MPMC_B200_DEVICES", "MPMC_B200_GPUS"):
            e.pop(v, None)
        e["MPMC_B200_DEVICE"] = str(devices[k % len(devices)])
        # the chains share the host: split the cores between their shims' host threads (8 chains x 8 spinning OpenMP
        # threads on 16 cores cost each chain 3.6x, profiles/r02_multi_gpu.md) and let idle threads sleep
        e.setdefault("MPMC_B200_HOST_THREADS", str(host_threads_per_chain(chains)))
        e.setdefault("OMP_WAIT_POLICY", "PASSIVE")
        procs.append((k, d, subprocess.Popen([program, "input"], cwd=d, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                                             env=e), time.perf_counter()))
    results = []
    for k, d, p, ts in procs:
        try:
            so, se = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            p.kill()
            so, se = p.communicate()
        results.append({"chain": k, "device": devices[k % len(devices)], "seed": seeds[k % len(seeds)], "rc": p.returncode,
                        "dir": d, "wall_s": time.perf_counter() - ts, "averages": parse_averages(so),
                        "error": (so[-400:] + se[-400:]) if p.returncode else None})
    wall = time.perf_counter() - t0
    out = {"chains": results, "combined": combine([r["averages"] for r in results if r["rc"] == 0]), "wall_s": wall,
           "workdir": None if own else workdir}
    if own:
        shutil.rmtree(workdir, ignore_errors=True)
    return out


def combine(per_chain: List[Dict[str, Dict[str, float]]]) -> Dict[str, Dict[str, float]]:
    """Mean over the chains of every averaged observable, with the standard error of that mean from the spread of the
    chain means (what the root rank's sdom does with the ranks' values, average.c:224-349)."""
    out: Dict[str, Dict[str, float]] = {}
    if not per_chain:
        return out
    for key in per_chain[0]:
        vals = [c[key]["value"] for c in per_chain if key in c]
        n = len(vals)
        mean = sum(vals) / n
        var = sum((v - mean) ** 2 for v in vals) / (n - 1) if n > 1 else 0.0
        out[key] = {"value": mean, "error": (var / n) ** 0.5 if n > 1 else per_chain[0][key]["error"], "chains": n,
                    "unit": per_chain[0][key].get("unit", "")}
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("input")
    ap.add_argument("--chains", type=int, default=8)
    ap.add_argument("--devices", default=None, help="e.g. 0-7 or 0,2,4 (default 0..chains-1)")
    ap.add_argument("--seeds", default=None, help="comma-separated (default 1..chains)")
    ap.add_argument("--keep", default=None, help="directory for the chains' files (default: a temporary one, removed)")
    a = ap.parse_args(argv)
    seeds = [int(x) for x in a.seeds.split(",")] if a.seeds else None
    r = launch(a.input, a.chains, parse_devices(a.devices, a.chains), seeds, a.keep)
    json.dump(r, sys.stdout, indent=1)
    print()
    return 0 if all(c["rc"] == 0 for c in r["chains"]) else 1


if __name__ == "__main__":
    sys.exit(main())
