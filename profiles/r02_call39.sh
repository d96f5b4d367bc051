#!/bin/bash
# round 2, GPU call 39 (one GPU): K4 half-tile items (MPMC_B200_SYM_HSPLIT=2) on one GPU at mid sizes, where the grid is ~1.3 waves
mkdir -p gpurun_out
for W in h2_fluid_16000 mof_h2_21k h2_fluid_32000 mof_h2_42k; do
  for H in 1 2; do
    MPMC_B200_SYM_HSPLIT=$H timeout 200 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c39_${W}_h$H.json 2> gpurun_out/r02_c39_${W}_h$H.err
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c39_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c39_')[1][:22].ljust(22), round(d['ms_per_step'],3), 'K4/launch', round(d['roofline'].get('ms_per_launch') or 0,4), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
