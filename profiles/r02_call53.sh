#!/bin/bash
# round 2, GPU call 53 (one GPU): the edited-atom-set parity test; uVT through the drop-in (PCN-61 + 500 H2, 8,548 atoms) with the
# spatial orders carried over and rebuilt
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "small_edits" > gpurun_out/r02_c53_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c53_tests.log
tail -3 gpurun_out/r02_c53_tests.log
python - > gpurun_out/r02_c53_uvt_dropin.txt 2>&1 <<'PY'
import os, sys, json, gzip, base64, subprocess, tempfile
sys.path.insert(0, os.getcwd())
import bench
GOLD = os.path.join(os.getcwd(), "tests", "golden")
name = "traj_pcn61_uvt"
d = json.load(open(os.path.join(GOLD, name + ".json")))
def unpack(key):
    return gzip.decompress(base64.b64decode(d[key + "_gz_b64"])).decode() if key + "_gz_b64" in d else d[key if key != "final" else "final_pqr"]
for reuse in ("1", "0", "1", "0"):
    with tempfile.TemporaryDirectory(prefix="uvt_") as td:
        open(os.path.join(td, d["pqr_name"]), "w").write(unpack("pqr"))
        open(os.path.join(td, "input"), "w").write(d["input"])
        for fn, text in (d.get("extra_files") or {}).items():
            open(os.path.join(td, fn), "w").write(text)
        env = dict(os.environ, MPMC_B200_SHIM_TIMING="1", MPMC_B200_REUSE_ORDER=reuse)
        r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
        print("MPMC_B200_REUSE_ORDER=" + reuse, "rc", r.returncode, [l.split("SHIM_TIMING: ")[1][:60] for l in r.stderr.splitlines() if "per second" in l])
PY
cat gpurun_out/r02_c53_uvt_dropin.txt
