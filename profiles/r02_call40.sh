#!/bin/bash
# round 2, GPU call 40 (one GPU): K4 half-tile items on one GPU at the headline size and at 64k-atom fluid
mkdir -p gpurun_out
for W in mof_h2_63k h2_fluid_64000 h2_fluid_8000; do
  for H in 1 2; do
    MPMC_B200_SYM_HSPLIT=$H timeout 300 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c40_${W}_h$H.json 2> gpurun_out/r02_c40_${W}_h$H.err
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c40_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c40_')[1][:22].ljust(22), round(d['ms_per_step'],3), 'K4/launch', round(d['roofline'].get('ms_per_launch') or 0,4), 'npol', d['config']['n_polarizable'], (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
