#!/bin/bash
# round 2, GPU call 23 (one GPU): K1 without per-pair sqrt / divisions (rsqrt + r^2 band test, per-atom sqrt(eps) and 1/m): parity, then timing
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pins.py tests/test_gpu_incremental.py tests/test_gpu_shim.py -m gpu -q -x > gpurun_out/r02_c23_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c23_tests.log
tail -6 gpurun_out/r02_c23_tests.log
for W in co2_fluid_12k h2_fluid_16000 h2_fluid_64000 mof_h2_63k; do
  timeout 300 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c23_bench_$W.json 2> gpurun_out/r02_c23_bench_$W.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c23_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c23_bench_')[1][:18].ljust(18), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-500:])
PY
