#!/bin/bash
# round 2, GPU call 56 (one GPU): after the device-side subset gather and the cheaper nearest-survivor search: set_atoms cost, uVT
# walks, whole GPU suite, default bench
mkdir -p gpurun_out
MPMC_B200_REPACK_TIMING=1 timeout 600 python profiles/set_atoms_cost.py > gpurun_out/r02_c56_set_atoms.txt 2> gpurun_out/r02_c56_set_atoms.err
cat gpurun_out/r02_c56_set_atoms.txt; grep "n 64" gpurun_out/r02_c56_set_atoms.err | tail -2
for W in mof_h2_21k mof_h2_63k; do timeout 600 python profiles/uvt_edit_walk.py $W 60 > gpurun_out/r02_c56_walk_$W.txt 2>&1; tail -1 gpurun_out/r02_c56_walk_$W.txt; done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c56_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c56_tests.log
tail -3 gpurun_out/r02_c56_tests.log
timeout 600 python bench.py > gpurun_out/r02_c56_bench.json 2> gpurun_out/r02_c56_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c56_bench.json').read().strip().splitlines()[-1])
print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'].get('value'), 'frac', d['roofline'].get('frac'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))
print('mc dropin', d['config'].get('mc_steps_per_s_dropin'), 'resident', d['config'].get('resident_steps'))
PY
