#!/bin/bash
# round 2, GPU call 77 (one GPU): does the crossover rows / symmetric move below 4,096 with the finest items?
mkdir -p gpurun_out
for w in h2_fluid_5000 h2_fluid_6000 h2_fluid_6600; do
for m in default 0; do
  if [ $m = default ]; then unset MPMC_B200_SYM_MIN; else export MPMC_B200_SYM_MIN=$m; fi
  timeout 60 python bench.py --steps 20 --warmup 5 --workload $w --no-cpu-baseline --no-mc-dropin 2> gpurun_out/r02_c77.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w SYM_MIN=$m', 'ms', round(d['ms_per_step'],4), 'dipole', round(d['roofline']['phase_ms']['dipole'],4))"
done
done
