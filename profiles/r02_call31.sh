#!/bin/bash
# round 2, GPU call 31 (one GPU): K1 with the Feynman-Hibbs derivative sums collapsed: the whole GPU suite, then the FH workloads
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c31_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c31_tests.log
tail -4 gpurun_out/r02_c31_tests.log
for W in co2_fluid_12k h2_fluid_16000 h2_fluid_64000 mof_h2_63k; do
  timeout 300 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c31_bench_$W.json 2> gpurun_out/r02_c31_bench_$W.err
done
MPMC_B200_OVERLAP=0 timeout 300 python bench.py --steps 10 --warmup 3 --workload h2_fluid_64000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c31_bench_h2_fluid_64000_nooverlap.json 2> gpurun_out/r02_c31_bench_h2_fluid_64000_nooverlap.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c31_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c31_bench_')[1][:30].ljust(30), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
