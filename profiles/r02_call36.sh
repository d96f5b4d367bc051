#!/bin/bash
# round 2, GPU call 36 (one GPU): the state after the K1 work -- whole GPU suite (-x), smoke, default bench line, and the drop-in MC
# steps/s of BASELINE config 2 (co2_fluid_12k) on the per-step path, with resident steps, and with the incremental evaluation
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c36_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c36_tests.log
tail -3 gpurun_out/r02_c36_tests.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02_c36_smoke.log 2>&1; tail -1 gpurun_out/r02_c36_smoke.log
timeout 600 python bench.py > gpurun_out/r02_c36_bench.json 2> gpurun_out/r02_c36_bench.err; echo "bench rc=$?"
RESIDENT_AB_ONLY=co2_fluid_12k timeout 600 python profiles/resident_steps_ab.py 0 64 > gpurun_out/r02_c36_co2_12k_dropin.txt 2>&1
RESIDENT_AB_ONLY=co2_fluid_12k MPMC_B200_INCREMENTAL=1 timeout 600 python profiles/resident_steps_ab.py 0 >> gpurun_out/r02_c36_co2_12k_dropin.txt 2>&1
cat gpurun_out/r02_c36_co2_12k_dropin.txt
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c36_bench.json').read().strip().splitlines()[-1])
print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'].get('value'), 'frac', d['roofline'].get('frac'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))
print('mc dropin', d['config'].get('mc_steps_per_s_dropin'), 'resident', d['config'].get('resident_steps'))
print('cpu', (d.get('cpu_baseline') or {}).get('value'))
PY
