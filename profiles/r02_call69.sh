#!/bin/bash
# round 2, GPU call 69 (one GPU): where a uVT step of PCN-61 + 500 H2 goes (shim's own clock, every phase line)
mkdir -p gpurun_out
python - > gpurun_out/r02_c69_uvt_phases.txt 2>&1 <<'PY'
import os, sys, json, gzip, base64, subprocess, tempfile
sys.path.insert(0, os.getcwd())
import bench
GOLD = os.path.join(os.getcwd(), "tests", "golden")
d = json.load(open(os.path.join(GOLD, "traj_pcn61_uvt.json")))
def unpack(key):
    return gzip.decompress(base64.b64decode(d[key + "_gz_b64"])).decode() if key + "_gz_b64" in d else d[key if key != "final" else "final_pqr"]
with tempfile.TemporaryDirectory(prefix="uvt_") as td:
    open(os.path.join(td, d["pqr_name"]), "w").write(unpack("pqr"))
    open(os.path.join(td, "input"), "w").write(d["input"])
    for fn, text in (d.get("extra_files") or {}).items():
        open(os.path.join(td, fn), "w").write(text)
    env = dict(os.environ, MPMC_B200_SHIM_TIMING="1", MPMC_B200_REPACK_TIMING="1")
    r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
    print("rc", r.returncode)
    lines = [l for l in r.stderr.splitlines() if "TIMING" in l]
    print("\n".join(lines[-60:]))
    print(d["input"])
PY
tail -80 gpurun_out/r02_c69_uvt_phases.txt
