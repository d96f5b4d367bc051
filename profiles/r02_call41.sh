#!/bin/bash
# round 2, GPU call 41 (one GPU): K4 half-tile items as the default on one GPU: whole GPU suite, executed-flop capture of the
# new launch shape, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c41_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c41_tests.log
tail -3 gpurun_out/r02_c41_tests.log
python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c41_probe.log 2>&1 || { echo probe failed; tail -5 gpurun_out/r02_c41_probe.log; }
timeout 900 ncu --metrics $(python profiles/capture_k4_executed.py --metrics) --clock-control none -k regex:dipole_sym_kernel -s 10 -c 2 --csv --log-file gpurun_out/r02_k4_executed_hsplit2_mof_h2_63k.csv python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c41_ncu.log 2>&1; echo "ncu rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c41_bench.json 2> gpurun_out/r02_c41_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c41_bench.json').read().strip().splitlines()[-1])
print('ms_per_step', d['ms_per_step'], 'K4/launch', d['roofline'].get('ms_per_launch'), 'frac(old capture)', d['roofline'].get('frac'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))
PY
