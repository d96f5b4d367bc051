#!/bin/bash
# round 2, GPU call 73 (one GPU): K4 items cut into groups of j-clusters (gridDim.z 4 / 8) for launches of less than two waves:
# parity of the K4-heavy tests with the automatic choice, then A/B at mid sizes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pins.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r02_c73_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c73_tests.log
tail -4 gpurun_out/r02_c73_tests.log
for w in h2_fluid_8000 mof_h2_7k h2_fluid_12000 h2_fluid_16000 mof_h2_21k; do
for h in 2 4 8; do
  MPMC_B200_SYM_HSPLIT=$h timeout 300 python bench.py --steps 20 --warmup 5 --workload $w --no-cpu-baseline --no-mc-dropin 2> gpurun_out/r02_c73.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w HSPLIT=$h', 'ms', round(d['ms_per_step'],4), 'dipole', round(d['roofline']['phase_ms']['dipole'],4), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))"
done
done
