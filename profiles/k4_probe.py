#!/usr/bin/env python
"""profiles/k4_probe.py -- the smallest program that launches the dominant kernel on a bench workload: builds the
system, evaluates energy() twice through the C ABI and exits.  Target of the ncu captures (profiles/capture_k4_executed.py);
never a source of timing numbers.

    python profiles/k4_probe.py [workload]            (default mof_h2_63k)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mpmc_b200 import Engine  # noqa: E402


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "mof_h2_63k"
    s = bench.build_workload(workload)
    eng = Engine(0)
    eng.set_system(s)
    for _ in range(2):
        e = eng.energy()
    c = eng.counters()
    print(f"k4_probe {workload}: E={e.total!r} polar={e.polar!r} iterations={e.polar_iterations} "
          f"launches={c.kernel_launches} contract_ms={eng.phase_ms('contract'):.4f}")
    eng.close()


if __name__ == "__main__":
    main()
