#!/bin/bash
# round 2, GPU call 65 (one GPU): K4 with one image per cluster pair as the default (damped classes at unroll 2): whole GPU
# suite, executed-flop capture of the new kernel, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c65_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c65_tests.log
tail -3 gpurun_out/r02_c65_tests.log
python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c65_probe.log 2>&1 || { echo probe failed; tail -5 gpurun_out/r02_c65_probe.log; }
timeout 900 ncu --metrics $(python profiles/capture_k4_executed.py --metrics) --clock-control none -k regex:dipole_sym_kernel -s 10 -c 2 --csv --log-file gpurun_out/r02_k4_executed_one_image_mof_h2_63k.csv python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c65_ncu.log 2>&1; echo "ncu rc=$?"
timeout 600 python bench.py > gpurun_out/r02_c65_bench.json 2> gpurun_out/r02_c65_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c65_bench.json').read().strip().splitlines()[-1])
print('ms_per_step', d['ms_per_step'], 'K4/launch', d['roofline'].get('ms_per_launch'), 'frac(old capture)', d['roofline'].get('frac'), 'e2e', d['e2e'].get('value'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))
PY
