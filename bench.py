#!/usr/bin/env python
"""bench.py -- the per-MC-step energy hot path (LJ + Ewald + iterative Thole) on B200.

One "step" = one full energy() evaluation of the workload (what one MC trial move costs in the
reference: src/mc/mc.c:302 calls energy() once per step and energy() recomputes everything).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

Own arm (default): the CUDA engine behind the C ABI (mpmc_b200/lib/libmpmc_b200.so).
  value  : Gpair-interactions/s, device-timed (CUDA events), system resident in HBM; one rank per GPU under torchrun,
           every exchange of a sharded evaluation done by the engine's kernels over NVLink peer memory
  e2e    : the same metric through the DROP-IN boundary: energy(system_t*) of the reference's own data model
           (mpmc_b200/shim: list pass, upload of what moved, evaluation on N GPUs from one process, write-back of
           mu / fields into every atom_t), timed by the same loop that times the reference in the CPU arm
  e2e_abi: the same through the C ABI with host buffers (ctypes), every rank copying positions in and arrays out
  config.parity : this run's energies and per-atom arrays against the pin of the workload (tests/golden/pin_*.npz:
           the unmodified reference where it fits, else the reference-pinned oracle), at every N
  config.mc_steps_per_s_dropin : MC steps/s of a short `mpmc_b200 <input>` NVT run (BASELINE's first metric)
  config.resident_steps : SURVEY 8f f3 -- MC steps/s of the drop-in on a small system, per-step path vs resident MC steps
  roofline: the dominant kernel (K4 symmetric dipole-tensor contraction): EXECUTED fp64 flops per launch (ncu SASS
           counts dadd + dmul + 2 dfma, profiles/k4_executed.json) / mean launch time (CUDA events on the launching
           stream) against the fp64 FMA peak of 148 SMs at the clock sampled under load; the SURVEY 8d algorithmic
           figure is reported beside it as `algorithmic_tflops`
  cpu_baseline: the unmodified reference (oracle/_ref) or the oracle port, on the host cores
Reference arm (--impl reference): the reference's own CPU energy() on the host cores.

The oracle/ directory is used ONLY for cpu_baseline and --impl reference (it is the thing timed
there, never a fallback for the CUDA path).  Nothing here reads /root/reference at run time.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "pair_interactions_per_second"
UNIT = "Gpair/s"

# algorithmic fp64 flops per unit of work
FLOP_GEOM, FLOP_PAIR_FH4, FLOP_PAIR, FLOP_FIELD, FLOP_DIPOLE, FLOP_RECIP = 51.0, 369.0, 153.0, 81.0, 126.0, 49.0


OPTION_SUFFIXES = ("_gs_production", "_ewald_full", "_es_wolf", "_rd_crystal1", "_rd_crystal2", "_rd_crystal3",
                   "_rd_crystal4")


def split_workload(name: str):
    """(base workload, option suffixes left to right): mof_h2_63k_ewald_full -> ("mof_h2_63k", ["_ewald_full"])."""
    found = []
    while True:
        for k in OPTION_SUFFIXES:
            if name.endswith(k):
                found.insert(0, k)
                name = name[:-len(k)]
                break
        else:
            return name, found


def apply_options(s, suffixes):
    """The option suffixes of a workload name applied to a system (the full-size one or the CPU arm's sample)."""
    p = s.params
    for k in suffixes:
        if k == "_gs_production":  # the solver options of MPMC's production inputs (Gauss-Seidel, single GPU)
            p.polar_gs_ranked, p.polar_max_iter, p.polar_palmo, p.polar_wolf, p.polar_wolf_alpha = 1, 4, 1, 1, 0.13
        elif k == "_ewald_full":  # the same system solved with full-Ewald polarization
            p.polar_ewald_full = 1
        elif k == "_es_wolf":  # coulombic_wolf() instead of the Ewald sum (the reference needs FH off)
            p.es_wolf, p.feynman_hibbs = 1, 0
        elif k.startswith("_rd_crystal"):  # LJ summed over the (2*order-1)^3 periodic images
            p.rd_crystal, p.rd_crystal_order = 1, int(k[-1])
    return s


def build_workload(name: str):
    """Synthetic systems of BASELINE.json's configs, optionally with option suffixes."""
    from mpmc_b200 import systems

    base, suffixes = split_workload(name)
    if suffixes:
        return apply_options(build_workload(base), suffixes)
    if name == "mof_h2_63k":  # config 4: MOF supercell + H2, ~63k atoms, full polarizable energy, Jacobi x10
        return systems.mof_h2()
    if name == "mof_h2_21k":  # config 3 shape: 128.388 x 42.796 x 42.796 A, Jacobi x4 (3_PCN61/iter.inp)
        return systems.mof_h2(cells=(9, 3, 3), polar_max_iter=4)
    if name == "mof_h2_42k":
        return systems.mof_h2(cells=(9, 6, 3))
    if name == "mof_h2_7k":
        return systems.mof_h2(cells=(3, 3, 3))
    if name == "co2_fluid_12k":  # config 2
        return systems.co2_fluid(4096)
    if name.startswith("h2_fluid_"):  # config 5 sweep
        return systems.h2_fluid(int(name.split("_")[-1]))
    raise SystemExit(f"unknown workload {name}")


def pair_interactions(cnt) -> float:
    """the pair terms the reference evaluates in one energy() call."""
    return (cnt.pairs_lj + cnt.pairs_es + cnt.pairs_field + 0.5 * cnt.pairs_dipole * cnt.contractions)


def algorithmic_flops(cnt, fh: bool) -> float:
    per_cut = FLOP_PAIR_FH4 if fh else FLOP_PAIR
    return (FLOP_GEOM * cnt.pairs_tested + per_cut * max(cnt.pairs_lj, cnt.pairs_es) + FLOP_FIELD * cnt.pairs_field +
            FLOP_DIPOLE * cnt.pairs_dipole * cnt.contractions + FLOP_RECIP * cnt.kvectors * cnt.recip_atoms)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU side: the reference's own implementation (oracle/_ref) or the oracle port, on the host cores
# ------------------------------------------------------------------------------------------------
def _cpu_sample_system(workload: str, budget_s: float, calls: int):
    """A bounded sample of the workload: the same generator and options on a smaller supercell, so
    that `calls` evaluations of the O(N^2) CPU path fit in `budget_s` (2.1e-7 N^2 s per reference
    call on one core of this class of host, from the 1,236-atom sample: 0.32 s)."""
    from mpmc_b200 import systems

    workload, suffixes = split_workload(workload)
    for k in suffixes:  # the image sums multiply the reference's LJ loop
        if k.startswith("_rd_crystal"):
            budget_s /= 1.0 + 0.3 * (2 * int(k[-1]) - 1) ** 3
    s, what = _cpu_sample_base(workload, budget_s, calls)
    if suffixes:
        what += " with the options of " + "".join(suffixes)
    return apply_options(s, suffixes), what


def _cpu_sample_base(workload: str, budget_s: float, calls: int):
    from mpmc_b200 import systems

    if workload.startswith("mof_h2"):
        for cells in ((3, 3, 3), (3, 3, 2), (2, 2, 2), (2, 2, 1)):
            n = int(round(63549 * cells[0] * cells[1] * cells[2] / 243.0))
            if 2.1e-7 * n * n * calls <= budget_s or cells == (2, 2, 1):
                return systems.mof_h2(cells=cells), f"mof_h2 supercell {cells[0]}x{cells[1]}x{cells[2]} of the 9x9x3 workload"
    if workload.startswith("co2_fluid"):
        for m in (1728, 1000, 512):
            if 0.5e-7 * (3 * m) ** 2 * calls <= budget_s or m == 512:
                return systems.co2_fluid(m), f"co2_fluid with {m} molecules"
    for n in (4000, 2000, 1000):
        if 2.1e-7 * n * n * calls <= budget_s or n == 1000:
            return systems.h2_fluid(n), f"h2_fluid with {n} atoms"


def _sample_pairs(sample) -> float:
    """Pair interactions of one energy() call on the sample (oracle counters; untimed bookkeeping)."""
    from oracle import oracle

    _, cnt, _ = oracle.energy(sample, nthreads=0)
    return pair_interactions(cnt), cnt


def run_cpu_reference(workload: str, steps: int, warmup: int, budget_s: float = 120.0):
    """Times energy() of the reference on the host cores.  With oracle/_ref present: the UNMODIFIED
    reference (single-threaded C, as shipped) as `cores` independent replicas, one per host core --
    the reference's own parallel mode (MPI replicas, src/mc/mc.c:431).  Otherwise the OpenMP oracle
    port on all cores.  Returns dict(value Gpair/s, cores, kind, sample, ms_per_step)."""
    from mpmc_b200 import systems
    from oracle import oracle

    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    sample, what = _cpu_sample_system(workload, budget_s, steps + warmup)
    pairs, cnt = _sample_pairs(sample)
    if os.path.exists(oracle.REF_DRIVER):
        # memory of the reference: 200 B per pair_t + 72 N^2 B for the A matrix
        per_proc = sample.n * sample.n * (100.0 + 72.0) + 1e9
        try:
            avail = os.sysconf("SC_AVPHYS_PAGES") * os.sysconf("SC_PAGE_SIZE")
        except Exception:
            avail = 64e9
        nproc = int(max(1, min(cores, 0.7 * avail // per_proc)))
        with tempfile.TemporaryDirectory(prefix="mpmc_ref_") as tmp:
            procs = []
            for r in range(nproc):
                wd = os.path.join(tmp, f"r{r}")
                inp = systems.write_reference_case(sample, wd)
                procs.append(subprocess.Popen([oracle.REF_DRIVER, os.path.basename(inp), str(steps), str(warmup)], cwd=wd,
                                              stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True))
            secs = []
            for p in procs:
                out, _ = p.communicate()
                t = [l for l in out.splitlines() if l.startswith("timing seconds_per_energy")]
                if p.returncode != 0 or not t:
                    raise RuntimeError("ref_driver failed:\n" + out[-2000:])
                secs.append(float(t[-1].split()[2]))
        sec = max(secs)  # replicas run concurrently: the slowest one bounds the step
        return dict(value=nproc * pairs / sec / 1e9, cores=nproc, kind="reference", ms_per_step=sec * 1e3,
                    sample=f"{what} (N={sample.n}, {pairs:.4g} pair interactions per energy()); {nproc} concurrent "
                           f"single-threaded replicas of the unmodified reference, {steps} timed energy() calls each "
                           f"after {warmup} warm-up, one molecule displaced between calls",
                    n_atoms=sample.n)
    nthreads = oracle.max_threads()
    for _ in range(warmup):
        oracle.energy(sample, nthreads=0)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.energy(sample, nthreads=0)
    sec = (time.perf_counter() - t0) / max(steps, 1)
    return dict(value=pairs / sec / 1e9, cores=nthreads, kind="port", ms_per_step=sec * 1e3,
                sample=f"{what} (N={sample.n}, {pairs:.4g} pair interactions per energy()); OpenMP oracle port "
                       f"on {nthreads} threads, {steps} timed calls after {warmup} warm-up", n_atoms=sample.n)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    r = run_cpu_reference(args.workload, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "cpu_sample": r["sample"]},
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
# own arm
# ------------------------------------------------------------------------------------------------
SHIM_BUILD = os.path.join(ROOT, "mpmc_b200", "shim", "_build")
SHIM_DRIVER = os.path.join(SHIM_BUILD, "shim_driver")   # energy(system_t*) timing loop over the drop-in library
DROPIN = os.path.join(SHIM_BUILD, "mpmc_b200")          # the drop-in program: the reference's MC driver over the CUDA engine


def dropin_env(n_gpus: int) -> dict:
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    if n_gpus > 1:
        env["MPMC_B200_DEVICES"] = ",".join(str(k) for k in range(n_gpus))  # one process, n GPUs: mpmc_b200_create_group
    return env


def dropin_energy_seconds(s, n_gpus: int, steps: int, warmup: int):
    """Seconds per energy(system_t*) call through the drop-in boundary: shim_driver (oracle/ref_driver.c's timing loop
    linked against the drop-in library -- the same loop that times the reference in the CPU arm) parses the system,
    then displaces one molecule and calls energy() `steps` times after `warmup` untimed calls.  Each call walks the
    reference's molecule / atom lists, uploads what changed, evaluates on `n_gpus` GPUs and writes mu / E_static /
    E_induced back into every atom_t.  Returns (seconds per call, error text)."""
    if not os.path.exists(SHIM_DRIVER):
        return None, f"{SHIM_DRIVER} missing (built where /root/reference is present)"
    from mpmc_b200 import systems

    with tempfile.TemporaryDirectory(prefix="mpmc_dropin_") as td:
        inp = systems.write_reference_case(s, td)
        r = subprocess.run([SHIM_DRIVER, os.path.basename(inp), str(max(steps, 2)), str(max(warmup, 1)), "q"], cwd=td,
                           capture_output=True, text=True, timeout=1200, env=dropin_env(n_gpus))
    t = [l for l in r.stdout.splitlines() if l.startswith("timing seconds_per_energy")]
    if r.returncode != 0 or not t:
        return None, (r.stdout[-400:] + r.stderr[-400:]).replace("\n", " | ")
    return float(t[-1].split()[2]), ""


def dropin_mc_steps_per_s(s, n_gpus: int, short: int, long_: int):
    """MC steps/s of `mpmc_b200 <input>` (NVT, fixed seed), read from the shim's own clock (MPMC_B200_SHIM_TIMING: energy()
    calls per second of wall clock between the 1st and the last call of ONE run of `long_` steps -- parsing, CUDA start-up and
    the first allocations lie before the end of the 1st call -- the initial energy; everything the unmodified driver does per step is included).  The earlier
    form, the wall-clock difference of a short and a long run, carried the start-up jitter of two fresh processes
    (+-0.5 s against ~3 s of stepping)."""
    import re

    if not os.path.exists(DROPIN):
        return None
    from mpmc_b200 import systems

    steps = max(long_, short + 3)
    with tempfile.TemporaryDirectory(prefix="mpmc_mc_") as td:
        ex = f"numsteps {steps}\ncorrtime {steps}\npreset_seeds 12345\nmove_factor 0.01\nrot_factor 0.01\n"
        systems.write_reference_case(s, td, ensemble="nvt", extra=ex)
        env = dict(dropin_env(n_gpus), MPMC_B200_SHIM_TIMING="1")
        r = subprocess.run([DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
    m = re.search(r"([\d.]+) energy\(\) calls per second", r.stderr)
    return float(m.group(1)) if (m and r.returncode == 0) else None


def dropin_resident_steps(base: str = "h2_fluid_1000", steps: int = 6000, batch: int = 64):
    """SURVEY 8f f3 evidence in the driver-run line: MC steps/s of `mpmc_b200 <input>` on a SMALL system (where the host
    round trip per step costs more than the kernels) on the per-step path and with resident MC steps
    (MPMC_B200_RESIDENT_STEPS: batches of steps moved, evaluated and decided on the device, the step replayed as a CUDA
    graph), both read from the shim's own clock between the 1st and the last energy() call (MPMC_B200_SHIM_TIMING)."""
    import re

    if not os.path.exists(DROPIN):
        return None
    from mpmc_b200 import systems

    s = build_workload(base)
    out = {"workload": base, "ensemble": "nvt", "steps": steps, "batch": batch}
    for key, b in (("mc_steps_per_s", 0), ("mc_steps_per_s_resident", batch)):
        with tempfile.TemporaryDirectory(prefix="mpmc_res_") as td:
            ex = f"numsteps {steps}\ncorrtime {steps}\npreset_seeds 12345\nmove_factor 0.01\nrot_factor 0.01\n"
            systems.write_reference_case(s, td, ensemble="nvt", extra=ex)
            env = dict(dropin_env(1), MPMC_B200_SHIM_TIMING="1", MPMC_B200_RESIDENT_STEPS=str(b))
            r = subprocess.run([DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=600, env=env)
        m = re.search(r"([\d.]+) energy\(\) calls per second", r.stderr)
        out[key] = float(m.group(1)) if (m and r.returncode == 0) else None
        if b:
            m = re.search(r"RESIDENT_STEPS: (\d+) batches, (\d+) steps answered from a batch, (\d+) batches cut short", r.stderr)
            out["batches_served_cut"] = [int(x) for x in m.groups()] if m else None
    return out


def executed_flops(workload: str, n_gpus: int):
    """fp64 flops ONE launch of the dominant kernel executes on one rank: dadd + dmul + 2 dfma thread-level SASS counts
    of an ncu capture of this workload (profiles/k4_executed.json, written by profiles/capture_k4_executed.py from the
    raw CSV committed beside it).  The work items of the symmetric contraction are dealt out evenly, so a rank of an
    n-GPU run executes 1/n of the single-GPU launch."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "k4_executed.json")))[workload]
    except Exception:
        return None
    rec = dict(rec)
    rec["flops_per_rank_launch"] = (rec["dadd"] + rec["dmul"] + 2.0 * rec["dfma"]) / max(n_gpus, 1)
    return rec


def own_arm(args):
    import torch
    import torch.distributed as dist

    from mpmc_b200 import Engine
    from tests import pins
    from mpmc_b200.sharded import ShardedEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the engine has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    s = build_workload(args.workload)
    fh = bool(s.params.feynman_hibbs)
    polar = bool(s.params.polarization and not s.params.rd_only)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    eng = Engine(local, stream=stream.cuda_stream)
    sh = ShardedEngine(eng, rank, world, stream=stream)
    sh.set_system(s)
    eng_peer = [bool(world > 1 and eng.peer_active())]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    flush = torch.empty(256 * 1024 * 1024 // 8, dtype=torch.float64, device="cuda")  # > 126 MB L2

    # ---- warm-up (untimed) ----
    for _ in range(max(args.warmup, 0)):
        sh.energy()
    peak_probe = eng.measure_fp64_peak()  # DFMA-chain probe on this GPU, same run (also warms the clocks)

    # ---- value: device-timed, system resident in HBM ----
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches = 0
    contract_ms = []
    phases = {}
    e = None
    for k in range(args.steps):
        flush.fill_(float(k))  # L2 flush between timed iterations (untimed)
        ev[k][0].record(stream)
        e = sh.energy()
        ev[k][1].record(stream)
        launches += sh.kernel_launches()
        contract_ms.append(eng.phase_ms("contract"))
        for ph in ("pair", "recip", "field", "dipole", "total"):
            phases.setdefault(ph, []).append(eng.phase_ms(ph))
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_dev = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([ms_dev], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    cnt = sh.counters_total()
    pairs = pair_interactions(cnt)
    value = pairs / (ms_per_step * 1e-3) / 1e9
    energy_timed = e.total

    # ---- parity of THIS run (the resident, unperturbed system) against the pin of the workload ----
    parity = None
    arr = sh.atom_arrays() if polar else None  # the caller's write-back buffers, reused below
    pin_name = pins.for_workload(args.workload)
    if pin_name is not None and rank == 0:
        pin = pins.load(pin_name)
        changed = pins.check_inputs(pin, s)
        dev = pins.deviation(pin, e.to_dict(), arr)
        parity = {"pin": pin_name, "pinned_by": pin["meta"]["pinned_by"], "inputs_match": not changed,
                  "max_rel_dev": dev["max"], "energies": dev["energies"], "arrays": dev["arrays"],
                  "checksums": dev["checksums"], "iterations_equal": dev["iterations_equal"], "worst": dev["worst"],
                  "ok": bool(not changed and dev["iterations_equal"] and dev["max"] <= 1e-10)}

    # ---- the same metric through the C ABI with HOST buffers (every rank uploads the positions the caller changed and
    # reads energies + per-atom arrays back); kept beside the drop-in figure below as `e2e_abi` ----
    n = s.n
    hpos = torch.from_numpy(s.pos.copy()).pin_memory()
    hnp = hpos.numpy()
    rng = np.random.Generator(np.random.PCG64(7))
    movable = np.flatnonzero(s.frozen == 0)
    sites = 5 if args.workload.startswith(("mof_h2", "h2_fluid")) else 3
    barrier()
    t_set = t_eval = t_wb = 0.0
    t0 = time.perf_counter()
    for k in range(args.steps):
        # one MC translation of one molecule, applied on the host copy the caller owns
        first = int(movable[0]) + sites * int(rng.integers(0, max(1, movable.size // sites)))
        hnp[first:first + sites] += rng.uniform(-0.05, 0.05, size=3)
        ta = time.perf_counter()
        sh.set_positions(0, hnp)           # host -> device, all positions
        tb = time.perf_counter()
        sh.energy()                        # device -> host: energies
        tc = time.perf_counter()
        if polar:
            arr = sh.atom_arrays(arr)      # device -> host: mu, ef_static, ef_induced, ... as polar() writes back
        td = time.perf_counter()
        t_set, t_eval, t_wb = t_set + (tb - ta), t_eval + (tc - tb), t_wb + (td - tc)
        launches += sh.kernel_launches()
    barrier()
    sec = time.perf_counter() - t0
    t = torch.tensor([sec], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    abi_sec_per_step = float(t.item()) / args.steps
    own = sh.counters_own()
    if world > 1:
        dist.destroy_process_group()  # no NCCL kernel may sit on the other GPUs while the drop-in process uses them
    if rank != 0:
        eng.close()
        return 0
    eng.close()
    del flush
    torch.cuda.empty_cache()

    # ---- e2e: through the drop-in boundary, energy(system_t*) of the reference's own data model ----
    dsec, derr = dropin_energy_seconds(s, world, args.steps, args.warmup)
    h2d = sites * 3 * 8                                     # the shim uploads the atoms that moved
    d2h = 32 * 8 + ((4 * 3 + 2) * 8 * n if polar else 0)    # scalars + what polar() leaves in every atom_t
    if dsec is not None:
        e2e = {"value": pairs / dsec / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "through": "energy(system_t*) of the drop-in library (shim_driver: list pass + upload + evaluation on "
                          f"{world} GPU(s) + write-back into atom_t), one molecule displaced per call",
               "ms_per_call": dsec * 1e3, "calls_per_s": 1.0 / dsec}
    else:  # never silently substitute: say that the drop-in figure is missing and why
        e2e = {"value": None, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "error": derr}
    e2e_abi = {"value": pairs / abi_sec_per_step / 1e9, "unit": UNIT, "h2d_bytes_per_step": n * 3 * 8, "d2h_bytes_per_step": d2h,
               "through": "C ABI with host buffers (ctypes): set_positions(all) + energy + get_atom_arrays on every rank",
               "calls_per_s": 1.0 / abi_sec_per_step,
               "host_ms_per_step": {"set_positions": 1e3 * t_set / args.steps, "energy": 1e3 * t_eval / args.steps,
                                    "write_back": 1e3 * t_wb / args.steps}}
    mc_dropin = None
    if not args.no_mc_dropin:
        est = max(ms_per_step * 1e-3, 2e-3)  # size the long run for ~3 s of stepping
        short = 5
        mc_dropin = dropin_mc_steps_per_s(s, world, short, short + int(min(max(3.0 / est, 20), 4000)))
    resident = dropin_resident_steps() if (not args.no_mc_dropin and world == 1) else None

    # ---- roofline of the dominant kernel (K4): EXECUTED fp64 flops / launch time / fp64 peak ----
    c_ms = float(np.mean([x for x in contract_ms if x > 0])) if any(x > 0 for x in contract_ms) else 0.0
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    peak_nominal = 148 * 64 * 2 * sm_mhz * 1e6 / 1e12  # 64 fp64 FMA lanes / clk / SM at the clock sampled under load
    k4_alg = FLOP_DIPOLE * own.pairs_dipole_own  # SURVEY 8d convention: 126 flop per ordered pair
    ex = executed_flops(args.workload, world)
    roofline = {"bound": "fp64", "kernel": "dipole_sym_kernel (K4)", "unit": "TFLOP/s", "peak": peak_nominal,
                "peak_source": f"148 SM x 64 fp64 FMA lanes x 2 x {sm_mhz:.0f} MHz (clock sampled during the timed region; "
                               "MEASURED_PEAKS.json carries no fp64 figure)",
                "peak_dfma_probe": peak_probe, "ms_per_launch": c_ms,
                "algorithmic_tflops": (k4_alg / (c_ms * 1e-3) / 1e12) if c_ms > 0 else None,
                "algorithmic_note": "SURVEY 8d convention (126 flop per ORDERED pair as the reference's formula counts "
                                    "them); the symmetric kernel does the work once per unordered pair in ~75 executed "
                                    "flop, so this figure exceeds the hardware peak and is NOT the roofline fraction",
                "whole_eval_algorithmic_tflops": algorithmic_flops(cnt, fh) / (ms_per_step * 1e-3) / 1e12,
                "phase_ms": {k: float(np.mean(v)) for k, v in phases.items()}}
    if ex is not None and c_ms > 0:
        ach = ex["flops_per_rank_launch"] / (c_ms * 1e-3) / 1e12
        roofline.update({"achieved": ach, "frac": ach / peak_nominal, "flops_per_launch": ex["flops_per_rank_launch"],
                         "flops_source": f"executed dadd + dmul + 2 dfma of the ncu capture {ex.get('source', 'profiles/')}"
                                         + (f", 1/{world} of the single-GPU launch per rank" if world > 1 else ""),
                         "traffic": (ex.get("dram_bytes") / world) if ex.get("dram_bytes") else None,
                         "sm_pipe_fp64_pct": ex.get("pipe_fp64_pct")})
    else:
        roofline.update({"achieved": None, "frac": None, "traffic": None,
                         "flops_source": "no executed-flop capture for this workload (profiles/capture_k4_executed.py)"})
    list_pairs = float(getattr(cnt, "pairs_induced_list", 0.0))
    if s.params.polar_ewald_full and list_pairs > 0 and c_ms > 0:
        # full-Ewald polarization: the per-pass kernel between the phase marks is the pair-list apply (K7), HBM-bound:
        # 48-byte entries streamed once + 3 x 8 bytes of field written per polarizable atom (the mu_j gathers hit L2)
        nbytes = 48.0 * list_pairs + 24.0 * cnt.n_polarizable
        try:
            hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
            src = "MEASURED_PEAKS.json hbm_gbs (driver-measured copy bandwidth)"
        except Exception:
            hbm, src = 7700.0, "fallback: nominal HBM3e bandwidth of B200_PROFILING.md (MEASURED_PEAKS.json absent)"
        ach = nbytes / (c_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": "induced_list_apply_kernel (K7, per solver pass)", "achieved": ach, "peak": hbm,
                    "unit": "GB/s", "frac": ach / hbm, "traffic": None, "peak_source": src, "bytes_per_launch": nbytes,
                    "list_pairs": list_pairs, "ms_per_launch": c_ms,
                    "phase_ms": {k: float(np.mean(v)) for k, v in phases.items()}}

    cpu = None
    if not args.no_cpu_baseline:
        r = run_cpu_reference(args.workload, 2, 1, budget_s=25.0)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": args.workload, "n_atoms": n, "n_polarizable": int(cnt.n_polarizable),
                   "polar_iterations": int(e.polar_iterations), "pair_interactions_per_step": pairs,
                   "mc_steps_per_s": 1e3 / ms_per_step, "mc_steps_per_s_dropin": mc_dropin,
                   "resident_steps": resident,
                   "l2": "flushed between timed iterations (256 MiB fill)",
                   "parallelism": f"rows{world}" if world > 1 else "single", "energy_K": energy_timed, "parity": parity,
                   "exchange": ("engine kernels over NVLink peer memory (static-field rows, induced field of every pass, "
                                "final scalars); no NCCL call in energy()") if world > 1 and eng_peer[0] else
                               ("NCCL collectives between staged kernels" if world > 1 else None)},
        "e2e": e2e, "e2e_abi": e2e_abi,
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
    }
    print(json.dumps(line))
    return 0


def ensembles_arm(args):
    """--workload ensembles_<base>: `--gpus` independent fixed-seed NVT chains of <base>, one per GPU (BASELINE config 5's
    second half; the reference's own parallel mode, src/mc/mc.c:431-432), each the drop-in program `mpmc_b200 <input>`
    started by mpmc_b200/ensembles.py with MPMC_B200_DEVICE=k and seeds 1..N.  value = pair interactions per second
    summed over the chains (weak scaling: every GPU carries a whole chain); steps/s per chain from the wall-time
    difference of a long and a short run.  Under torchrun rank 0 launches all chains; the other ranks leave."""
    from mpmc_b200 import Engine, ensembles, systems

    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    base = args.workload[len("ensembles_"):]
    s = build_workload(base)
    eng = Engine(0)
    eng.set_system(s)
    for _ in range(max(args.warmup, 1)):
        e = eng.energy()
    cnt = eng.counters()
    pairs = pair_interactions(cnt)
    ms_one = eng.phase_ms("total")
    launches_per_step = cnt.kernel_launches
    eng.close()
    n = args.gpus
    est = max(ms_one * 1e-3 * 1.3, 1e-3)
    short, long_ = 5, 5 + int(min(max(4.0 / est, 20 * args.steps), 20000))
    sampler = ClockSampler(0)
    walls = {}
    with tempfile.TemporaryDirectory(prefix="mpmc_ens_") as td:
        for steps in (short, long_):
            case = os.path.join(td, f"case_{steps}")
            ex = f"numsteps {steps}\ncorrtime {steps}\nmove_factor 0.01\nrot_factor 0.01\n"
            systems.write_reference_case(s, case, ensemble="nvt", extra=ex)
            if steps == long_:
                sampler.start()
            r = ensembles.launch(os.path.join(case, "input"), n, devices=list(range(n)), seeds=list(range(1, n + 1)),
                                 workdir=os.path.join(td, f"run_{steps}"), env=dropin_env(1))
            bad = [c for c in r["chains"] if c["rc"] != 0]
            if bad:
                raise SystemExit(f"ensemble chain failed: {bad[0]['error']}")
            walls[steps] = [c["wall_s"] for c in r["chains"]]
            averages = r["combined"]
    clocks = sampler.stop()
    rates = [(long_ - short) / max(b - a, 1e-9) for a, b in zip(walls[short], walls[long_])]
    value = sum(rates) * pairs / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n, "steps": long_ - short, "warmup": short,
        "ms_per_step": 1e3 / min(rates), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": args.workload, "n_atoms": s.n, "chains": n, "seeds": list(range(1, n + 1)),
                   "pair_interactions_per_step": pairs, "mc_steps_per_s_per_chain": rates, "mc_steps_per_s": sum(rates),
                   "parallelism": f"ensembles{n} (independent chains, no collective)", "l2": "system larger than L2 or "
                   "flushed by the MC driver's own host work between steps",
                   "ensemble_averages": {k: averages[k] for k in ("potential energy", "N", "polarization energy") if k in averages}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 15 * 8, "d2h_bytes_per_step": 32 * 8 + 14 * 8 * s.n,
                "through": "mpmc_b200 <input>: the reference's MC driver, every step through energy(system_t*)"},
        "gpu_launches": int(launches_per_step * sum(long_ for _ in range(n))), "clocks": clocks, "roofline": None, "cpu_baseline": None,
    }
    if not args.no_cpu_baseline:
        r = run_cpu_reference(base, 2, 1, budget_s=25.0)
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="mof_h2_63k")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-mc-dropin", action="store_true", help="skip the short `mpmc_b200 <input>` MC run")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.workload.startswith("ensembles_"):
            args.workload = args.workload[len("ensembles_"):]  # N concurrent replicas of the reference is what that arm times
        return reference_arm(args)
    if args.workload.startswith("ensembles_"):
        return ensembles_arm(args)
    return own_arm(args)


if __name__ == "__main__":
    sys.exit(main())
