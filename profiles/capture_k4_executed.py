#!/usr/bin/env python
"""profiles/capture_k4_executed.py -- executed fp64 flops, DRAM traffic and fp64-pipe utilisation of ONE launch of the
dominant kernel, from an ncu metrics capture; the source of bench.py's `roofline.frac`.

On the GPU box (one GPU; after `python profiles/k4_probe.py <workload>` has exited 0):

    ncu --metrics $(python profiles/capture_k4_executed.py --metrics) --clock-control none \
        -k regex:dipole_sym_kernel -s 10 -c 2 --csv --log-file gpurun_out/r02_k4_executed_<workload>.csv \
        python profiles/k4_probe.py <workload>

Here (no GPU needed): turn the raw CSV into the record bench.py reads and keep the CSV beside it:

    python profiles/capture_k4_executed.py <workload> gpurun_out/r02_k4_executed_<workload>.csv

executed flops = dadd + dmul + 2 dfma, thread-level, predicated-on SASS counts of the launch (`-s 10` skips the first
evaluation's ten launches, so the launch counted runs with warm instruction caches; counts do not depend on that).
"""
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = {
    "dadd": "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
    "dmul": "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
    "dfma": "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum",
    "dram_read": "dram__bytes_read.sum",
    "dram_write": "dram__bytes_write.sum",
    "pipe_fp64_pct": "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "duration_ns": "gpu__time_duration.sum",
    "inst_executed": "smsp__inst_executed.sum",
    "threads_per_inst": "smsp__thread_inst_executed_per_inst_executed.ratio",
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6,
              "second": 1e9}


def parse(path):
    """-> list of {metric key: value} per captured launch, in capture order"""
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    by_id = {}
    for r in rd:
        name = r.get("Metric Name")
        for key, m in METRICS.items():
            if name == m:
                v = float(r["Metric Value"].replace(",", ""))
                v *= UNIT_SCALE.get(r.get("Metric Unit", ""), 1.0)
                by_id.setdefault(r["ID"], {"kernel": r.get("Kernel Name", "")})[key] = v
    for k in sorted(by_id, key=lambda x: int(x)):
        rows.append(by_id[k])
    return rows


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--metrics":
        print(",".join(METRICS.values()))
        return
    workload, raw = sys.argv[1], sys.argv[2]
    rows = parse(raw)
    if not rows:
        raise SystemExit(f"no metric rows in {raw}")
    r = rows[-1]
    keep = os.path.join(ROOT, "profiles", f"r02_k4_executed_{workload}.csv")
    if os.path.abspath(raw) != keep:
        shutil.copy(raw, keep)
    out = os.path.join(ROOT, "profiles", "k4_executed.json")
    try:
        rec = json.load(open(out))
    except Exception:
        rec = {}
    flops = r["dadd"] + r["dmul"] + 2.0 * r["dfma"]
    rec[workload] = {"dadd": r["dadd"], "dmul": r["dmul"], "dfma": r["dfma"], "flops": flops,
                     "dram_bytes": r.get("dram_read", 0.0) + r.get("dram_write", 0.0),
                     "pipe_fp64_pct": r.get("pipe_fp64_pct"), "ncu_duration_ms": r.get("duration_ns", 0.0) / 1e6,
                     "kernel": r.get("kernel", ""), "launches_captured": len(rows), "n_gpus": 1,
                     "source": os.path.relpath(keep, ROOT)}
    json.dump(rec, open(out, "w"), indent=1, sort_keys=True)
    print(json.dumps(rec[workload], indent=1))


if __name__ == "__main__":
    main()
