#!/bin/bash
# round 2, GPU call 42 (one GPU): K4 offset slices per item (MPMC_B200_SYM_CTAS_PER_SM) re-scanned with half-tile items
mkdir -p gpurun_out
for P in 16 24 32 48 72 96; do
  MPMC_B200_SYM_CTAS_PER_SM=$P timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c42_p$P.json 2> gpurun_out/r02_c42_p$P.err
done
for P in 24 48 96; do
  MPMC_B200_SYM_CTAS_PER_SM=$P timeout 300 python bench.py --steps 10 --warmup 3 --workload h2_fluid_32000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c42_h32_p$P.json 2> gpurun_out/r02_c42_h32_p$P.err
done
python - <<'PY'
import json,glob,re
for f in sorted(glob.glob('gpurun_out/r02_c42_*.json'), key=lambda x:(('h32' in x), int(re.findall(r'_p(\d+)',x)[0]))):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c42_')[1][:14].ljust(14), round(d['ms_per_step'],3), 'K4/launch', round(d['roofline'].get('ms_per_launch') or 0,4))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
