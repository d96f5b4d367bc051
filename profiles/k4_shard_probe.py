#!/usr/bin/env python
"""profiles/k4_shard_probe.py -- what ONE rank of an N-GPU run launches per solver pass, timed on a single GPU: the
context is sharded (rank r of `world`) and driven through the staged ABI without any exchange (the fields are then
incomplete -- timing only, never a source of results).  Prints the mean time of the contraction launch group
(dipole_sym_kernel + partial-row reduce) per knob setting; under `ncu --metrics gpu__time_duration.sum` the launch list
splits it by kernel.

    python profiles/k4_shard_probe.py [world ...]        (default 8 4 2)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mpmc_b200 import Engine, api  # noqa: E402


def run(world, rank, env, passes=10):
    for k, v in env.items():
        os.environ[k] = v
    eng = Engine(0)
    eng.set_shard(rank, world)
    eng.set_system(S)
    out = []
    for rep in range(3):
        eng.energy_enqueue()
        eng.polar_stage(api.STAGE_FIELD)
        eng.polar_stage(api.STAGE_INIT)
        for it in range(1, passes + 1):
            eng.polar_stage(api.STAGE_PASS, it)
            eng.polar_stage(api.STAGE_UPDATE, it)
            eng.polar_stage(api.STAGE_COMMIT, it)
        eng.polar_stage(api.STAGE_ENERGY, passes)
        eng.energy_finish()
        out.append((eng.phase_ms("contract"), eng.phase_ms("dipole"), eng.phase_ms("field"), eng.phase_ms("pair"),
                    eng.phase_ms("total")))
    eng.close()
    for k in env:
        os.environ.pop(k, None)
    return out[-1]


if __name__ == "__main__":
    one = "--one" in sys.argv  # default knobs only, rank 0 (the ncu launch-list target)
    worlds = [int(x) for x in sys.argv[1:] if x.isdigit()] or [8, 4, 2]
    S = bench.build_workload("mof_h2_63k")
    if one:
        for w in worlds:
            c, d, f, p, t = run(w, 0, {})
            print(f"world {w} rank 0: contract {c:.4f} ms (ideal {3.656 / w:.4f}), dipole phase {d:.3f}, total {t:.3f}", flush=True)
        sys.exit(0)
    for w in worlds:
        for env in ({}, {"MPMC_B200_SYM_CTAS_PER_SM": "24"}, {"MPMC_B200_SYM_CTAS_PER_SM": "12"},
                    {"MPMC_B200_SYM_CTAS_PER_SM": "8"}, {"MPMC_B200_SYM_HSPLIT": "2"}):
            c, d, f, p, t = run(w, 0, dict(env))
            print(f"world {w} rank 0 {env}: contract {c:.4f} ms (ideal {3.656 / w:.4f}), dipole phase {d:.3f}, field {f:.3f}, "
                  f"pair {p:.3f}, total {t:.3f}", flush=True)
        c, d, f, p, t = run(w, w - 1, {})
        print(f"world {w} rank {w - 1}: contract {c:.4f} ms, dipole phase {d:.3f}, field {f:.3f}, pair {p:.3f}, total {t:.3f}", flush=True)
