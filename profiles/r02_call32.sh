#!/bin/bash
# round 2, GPU call 32 (one GPU): K4 far-cluster threshold l r >= 48 (bit-exact skip) against 44 / 40 / 36 (A/B, parity printed)
mkdir -p gpurun_out
for LR in 48 44 40 36; do
  MPMC_B200_SYM_FAR_LR=$LR timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c32_bench_lr$LR.json 2> gpurun_out/r02_c32_bench_lr$LR.err
done
MPMC_B200_SYM_FAR_LR=40 timeout 300 python bench.py --steps 10 --warmup 3 --workload h2_fluid_64000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c32_bench_fluid_lr40.json 2> gpurun_out/r02_c32_bench_fluid_lr40.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c32_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c32_bench_')[1][:18].ljust(18), round(d['ms_per_step'],3), 'K4 ms/launch', round(d['roofline'].get('ms_per_launch') or 0,4), 'parity', d['config'].get('parity'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
