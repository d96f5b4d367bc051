#!/bin/bash
# round 2, GPU call 35 (one GPU): K1 slice count, finer
mkdir -p gpurun_out
for S in 39 52 78 104 193; do
  MPMC_B200_K1_SLICES=$S timeout 200 python bench.py --steps 10 --warmup 3 --workload co2_fluid_12k --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c35_co2_s$S.json 2> gpurun_out/r02_c35_co2_s$S.err
done
python - <<'PY'
import json,glob,re
for f in sorted(glob.glob('gpurun_out/r02_c35_*.json'), key=lambda x:(x.split('_s')[0], int(re.findall(r'_s(\d+)',x)[0]))):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c35_')[1][:14].ljust(14), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
