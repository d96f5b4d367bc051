#!/bin/bash
# round 2, GPU call 66 (one GPU): where the symmetric contraction (half-tile items) overtakes the row form on one GPU:
# MPMC_B200_SYM_MIN default (8192 padded polarizable atoms) against 0 (always symmetric)
mkdir -p gpurun_out
for w in h2_fluid_4000 h2_fluid_6000 h2_fluid_8000 h2_fluid_10000 h2_fluid_12000 mof_h2_7k; do
for m in default 0; do
  if [ $m = default ]; then unset MPMC_B200_SYM_MIN; else export MPMC_B200_SYM_MIN=$m; fi
  timeout 300 python bench.py --steps 20 --warmup 5 --workload $w --no-cpu-baseline --no-mc-dropin 2> gpurun_out/r02_c66.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w SYM_MIN=$m', 'ms', round(d['ms_per_step'],4), 'dipole', round(d['roofline']['phase_ms']['dipole'],4), 'launches', d.get('gpu_launches'))"
done
done
