"""bench.py's roofline numerator (executed fp64 flops of one K4 launch, profiles/k4_executed.json) must be what the committed
ncu metrics capture says (profiles/r02_k4_executed_<workload>.csv, parsed by profiles/capture_k4_executed.py): the record is
derived, never written by hand.  CPU test: reads the two committed files only."""
import importlib.util
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _capture_module():
    spec = importlib.util.spec_from_file_location("capture_k4_executed", os.path.join(ROOT, "profiles", "capture_k4_executed.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_executed_flop_record_matches_the_committed_ncu_capture():
    cap = _capture_module()
    rec = json.load(open(os.path.join(ROOT, "profiles", "k4_executed.json")))
    assert "mof_h2_63k" in rec  # the workload of the default bench line
    for workload, r in rec.items():
        raw = os.path.join(ROOT, r["source"])
        assert os.path.exists(raw), raw
        rows = cap.parse(raw)
        assert rows and len(rows) == r["launches_captured"]
        last = rows[-1]
        assert "dipole_sym_kernel" in last["kernel"] and last["kernel"] == r["kernel"]
        for k in ("dadd", "dmul", "dfma"):
            assert last[k] == r[k], (workload, k)
        assert r["flops"] == last["dadd"] + last["dmul"] + 2.0 * last["dfma"]
        assert r["dram_bytes"] == last["dram_read"] + last["dram_write"]
        assert abs(r["ncu_duration_ms"] - last["duration_ns"] / 1e6) < 1e-9
        # sanity of the capture itself: both captured launches of the same kernel execute the same work
        assert all(abs(x["dfma"] - last["dfma"]) <= 1e-6 * last["dfma"] for x in rows)


def test_bench_reads_that_record():
    import bench

    ex = bench.executed_flops("mof_h2_63k", 1)
    rec = json.load(open(os.path.join(ROOT, "profiles", "k4_executed.json")))["mof_h2_63k"]
    assert ex["flops_per_rank_launch"] == rec["dadd"] + rec["dmul"] + 2.0 * rec["dfma"]
    assert bench.executed_flops("mof_h2_63k", 8)["flops_per_rank_launch"] == ex["flops_per_rank_launch"] / 8
    assert bench.executed_flops("no_such_workload", 1) is None
