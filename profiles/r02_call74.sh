#!/bin/bash
# round 2, GPU call 74 (one GPU): final state -- whole GPU suite, smoke, default bench line, PCN-61 uVT through the drop-in, two mid sizes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c74_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c74_tests.log
tail -3 gpurun_out/r02_c74_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/r02_c74_bench.json 2> gpurun_out/r02_c74_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c74_bench.json').read().strip().splitlines()[-1])
print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'K4/launch', d['roofline'].get('ms_per_launch'), 'frac', d['roofline'].get('frac'), 'e2e', d['e2e'].get('value'), 'mc', d['config'].get('mc_steps_per_s_dropin'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))
PY
for w in h2_fluid_8000 mof_h2_7k; do
  timeout 300 python bench.py --steps 20 --warmup 5 --workload $w --no-cpu-baseline --no-mc-dropin 2>> gpurun_out/r02_c74.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w auto', 'ms', round(d['ms_per_step'],4), 'dipole', round(d['roofline']['phase_ms']['dipole'],4))"
done
python - > gpurun_out/r02_c74_uvt_dropin.txt 2>&1 <<'PY'
import os, sys, json, gzip, base64, subprocess, tempfile
sys.path.insert(0, os.getcwd())
import bench
GOLD = os.path.join(os.getcwd(), "tests", "golden")
d = json.load(open(os.path.join(GOLD, "traj_pcn61_uvt.json")))
def unpack(key):
    return gzip.decompress(base64.b64decode(d[key + "_gz_b64"])).decode() if key + "_gz_b64" in d else d[key if key != "final" else "final_pqr"]
for rep in range(2):
    with tempfile.TemporaryDirectory(prefix="uvt_") as td:
        open(os.path.join(td, d["pqr_name"]), "w").write(unpack("pqr"))
        open(os.path.join(td, "input"), "w").write(d["input"])
        for fn, text in (d.get("extra_files") or {}).items():
            open(os.path.join(td, fn), "w").write(text)
        r = subprocess.run([bench.DROPIN, "input"], cwd=td, capture_output=True, text=True, timeout=1200, env=dict(os.environ, MPMC_B200_SHIM_TIMING="1"))
        print("run", rep, "rc", r.returncode, [l.split("SHIM_TIMING: ")[1][:60] for l in r.stderr.splitlines() if "per second" in l])
PY
cat gpurun_out/r02_c74_uvt_dropin.txt
