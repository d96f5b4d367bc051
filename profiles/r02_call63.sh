#!/bin/bash
# round 2, GPU call 63 (one GPU): per-class unroll factors of the one-image K4 (O = the kernel of the previous commit, same box)
mkdir -p gpurun_out
for v in O E F G H I O; do
  MPMC_B200_LIB=$PWD/mpmc_b200/lib/variants/lib_$v.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin 2> gpurun_out/r02_c63_$v.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('variant $v', round(d['ms_per_step'],3), round(d['roofline']['phase_ms']['dipole'],3), (d['config'].get('parity') or {}).get('max_rel_dev'))"
done
