#!/bin/bash
# round 2, GPU call 22 (one GPU): the whole GPU suite the way the driver runs it, smoke, the default bench line (both arms),
# then -- after the plain runs exited 0 -- the ncu launch list of the bench command and one full capture of K4
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c22_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c22_tests.log
tail -4 gpurun_out/r02_c22_tests.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02_c22_smoke.log 2>&1; tail -1 gpurun_out/r02_c22_smoke.log
timeout 600 python bench.py > gpurun_out/r02_c22_bench.json 2> gpurun_out/r02_c22_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_c22_bench_reference.json 2> gpurun_out/r02_c22_bench_reference.err; echo "reference arm rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches_bench_mof63k.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c22_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dipole_sym_kernel -s 12 -c 1 -o gpurun_out/r02_k4_final -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c22_ncu_k4.log 2>&1
ncu -i gpurun_out/r02_k4_final.ncu-rep --page details --csv > gpurun_out/r02_k4_final_ncu_details.csv 2>/dev/null
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c22_bench.json').read().strip().splitlines()[-1])
print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'].get('value'), 'frac', d['roofline'].get('frac'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))
print('mc dropin', d['config'].get('mc_steps_per_s_dropin'), 'resident', d['config'].get('resident_steps'))
print('cpu', d.get('cpu_baseline'))
r=json.loads(open('gpurun_out/r02_c22_bench_reference.json').read().strip().splitlines()[-1]); print('reference arm', r.get('value'), r.get('unit'), r.get('cpu_baseline',{}).get('cores'))
PY
