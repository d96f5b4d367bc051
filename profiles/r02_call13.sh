#!/bin/bash
# round 2, GPU call 13 (one GPU): K3 (rsqrt path, r^2 band test, 8 warps per i-cluster), K1 fine slices for small systems
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02_c13_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c13_tests.log
for W in h2_fluid_1000 h2_fluid_2000 h2_fluid_4000 h2_fluid_8000 h2_fluid_16000 h2_fluid_64000 mof_h2_7k mof_h2_21k mof_h2_63k co2_fluid_12k; do
  timeout 200 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c13_bench_$W.json 2> gpurun_out/r02_c13_bench_$W.err
done
tail -4 gpurun_out/r02_c13_tests.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c13_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('bench_')[1][:22].ljust(22), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e)
PY
