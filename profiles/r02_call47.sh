#!/bin/bash
# round 2, GPU call 47 (one GPU): where a uVT step goes through the drop-in (insert / remove change the atom set: set_atoms + repack):
# the shipped PCN-61 + 500 H2 uVT case (8,548 atoms) and the small uVT trajectories, MPMC_B200_SHIM_TIMING
mkdir -p gpurun_out
python - > gpurun_out/r02_c47_uvt_timing.txt 2>&1 <<'PY'
import os, sys, json, gzip, base64, subprocess, tempfile
sys.path.insert(0, os.getcwd())
import bench
GOLD = os.path.join(os.getcwd(), "tests", "golden")
for name in ("traj_pcn61_uvt", "traj_co2_uvt", "traj_co2_nvt"):
    d = json.load(open(os.path.join(GOLD, name + ".json")))
    def unpack(key):
        return gzip.decompress(base64.b64decode(d[key + "_gz_b64"])).decode() if key + "_gz_b64" in d else d[key if key != "final" else "final_pqr"]
    with tempfile.TemporaryDirectory(prefix="uvt_") as td:
        open(os.path.join(td, d["pqr_name"]), "w").write(unpack("pqr"))
        open(os.path.join(td, "input"), "w").write(d["input"])
        for fn, text in (d.get("extra_files") or {}).items():
            open(os.path.join(td, fn), "w").write(text)
        env = dict(os.environ, MPMC_B200_SHIM_TIMING="1")
        r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=env)
        print(name, "rc", r.returncode)
        for l in r.stderr.splitlines():
            if "SHIM_TIMING" in l:
                print("   ", l)
PY
cat gpurun_out/r02_c47_uvt_timing.txt
