#!/bin/bash
# round 2, GPU call 72 (one GPU): Gauss-Seidel with the production options at small and mid sizes (persistent sweep / per-block launches)
mkdir -p gpurun_out
for w in h2_fluid_2000_gs_production mof_h2_7k_gs_production mof_h2_21k_gs_production; do
for g in default blocks; do
  if [ $g = default ]; then unset MPMC_B200_GS; else export MPMC_B200_GS=$g; fi
  timeout 300 python bench.py --steps 10 --warmup 3 --workload $w --no-cpu-baseline --no-mc-dropin 2> gpurun_out/r02_c72.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w GS=$g', 'ms', round(d['ms_per_step'],4), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, 'launches', d.get('gpu_launches'))"
done
done
