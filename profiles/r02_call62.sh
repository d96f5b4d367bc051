#!/bin/bash
# round 2, GPU call 62 (one GPU): code-size variants of the one-image-per-cluster-pair K4 (B: unroll 2; C: run-time switch, unroll 4;
# D: run-time switch, unroll 2), each with the switch off and on, bench workload
mkdir -p gpurun_out
for v in B C D; do
for u in 0 1; do
  MPMC_B200_LIB=$PWD/mpmc_b200/lib/variants/lib_$v.so MPMC_B200_SYM_UNISHIFT=$u timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin 2> gpurun_out/r02_c62_$v$u.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('variant $v unishift $u', round(d['ms_per_step'],3), d['roofline']['phase_ms']['dipole'], (d['config'].get('parity') or {}).get('max_rel_dev'))"
done
done
