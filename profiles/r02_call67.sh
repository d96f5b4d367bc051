#!/bin/bash
# round 2, GPU call 67 (one GPU): symmetric K4 from 4,096 padded polarizable atoms: whole GPU suite; PCN-61 + 500 H2 uVT through the
# drop-in with the old and the new threshold
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c67_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c67_tests.log
tail -3 gpurun_out/r02_c67_tests.log
python - > gpurun_out/r02_c67_uvt_dropin.txt 2>&1 <<'PY'
import os, sys, json, gzip, base64, subprocess, tempfile
sys.path.insert(0, os.getcwd())
import bench
GOLD = os.path.join(os.getcwd(), "tests", "golden")
name = "traj_pcn61_uvt"
d = json.load(open(os.path.join(GOLD, name + ".json")))
def unpack(key):
    return gzip.decompress(base64.b64decode(d[key + "_gz_b64"])).decode() if key + "_gz_b64" in d else d[key if key != "final" else "final_pqr"]
for symmin in ("8192", "4096", "8192", "4096"):
    with tempfile.TemporaryDirectory(prefix="uvt_") as td:
        open(os.path.join(td, d["pqr_name"]), "w").write(unpack("pqr"))
        open(os.path.join(td, "input"), "w").write(d["input"])
        for fn, text in (d.get("extra_files") or {}).items():
            open(os.path.join(td, fn), "w").write(text)
        env = dict(os.environ, MPMC_B200_SHIM_TIMING="1", MPMC_B200_SYM_MIN=symmin)
        r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
        print("MPMC_B200_SYM_MIN=" + symmin, "rc", r.returncode, [l.split("SHIM_TIMING: ")[1][:60] for l in r.stderr.splitlines() if "per second" in l])
PY
cat gpurun_out/r02_c67_uvt_dropin.txt
