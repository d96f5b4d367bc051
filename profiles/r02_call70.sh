#!/bin/bash
# round 2, GPU call 70 (one GPU): set_atoms reads a uVT step as two small edits (restore of the last trial + the new trial) instead of
# one long changed span: the new test, PCN-61 + 500 H2 uVT through the drop-in, whole GPU suite
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "two_edits or small_edits or large_moves" > gpurun_out/r02_c70_new_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c70_new_tests.log
tail -5 gpurun_out/r02_c70_new_tests.log
python - > gpurun_out/r02_c70_uvt_dropin.txt 2>&1 <<'PY'
import os, sys, json, gzip, base64, subprocess, tempfile
sys.path.insert(0, os.getcwd())
import bench
GOLD = os.path.join(os.getcwd(), "tests", "golden")
d = json.load(open(os.path.join(GOLD, "traj_pcn61_uvt.json")))
def unpack(key):
    return gzip.decompress(base64.b64decode(d[key + "_gz_b64"])).decode() if key + "_gz_b64" in d else d[key if key != "final" else "final_pqr"]
for rep in range(3):
    with tempfile.TemporaryDirectory(prefix="uvt_") as td:
        open(os.path.join(td, d["pqr_name"]), "w").write(unpack("pqr"))
        open(os.path.join(td, "input"), "w").write(d["input"])
        for fn, text in (d.get("extra_files") or {}).items():
            open(os.path.join(td, fn), "w").write(text)
        env = dict(os.environ, MPMC_B200_SHIM_TIMING="1")
        if rep == 2:
            env["MPMC_B200_REPACK_TIMING"] = "1"
        r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
        print("run", rep, "rc", r.returncode, [l.split("SHIM_TIMING: ")[1][:60] for l in r.stderr.splitlines() if "per second" in l])
        if rep == 2:
            ls = [l for l in r.stderr.splitlines() if "REPACK_TIMING" in l]
            print("repacks", len(ls), "carried", sum("carried" in l for l in ls), "full", sum("full sort" in l for l in ls))
PY
cat gpurun_out/r02_c70_uvt_dropin.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c70_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c70_tests.log
tail -3 gpurun_out/r02_c70_tests.log
