#!/bin/bash
# round 2, GPU call 34 (one GPU): K1 slice count (wave quantisation) A/B
mkdir -p gpurun_out
for S in 0 6 9 18 26 39; do
  MPMC_B200_K1_SLICES=$S timeout 200 python bench.py --steps 10 --warmup 3 --workload co2_fluid_12k --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c34_co2_s$S.json 2> gpurun_out/r02_c34_co2_s$S.err
done
for S in 0 4 6 12 16 24; do
  MPMC_B200_K1_SLICES=$S timeout 200 python bench.py --steps 10 --warmup 3 --workload h2_fluid_16000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c34_h2_s$S.json 2> gpurun_out/r02_c34_h2_s$S.err
done
for S in 0 2 4 6 9; do
  MPMC_B200_K1_SLICES=$S timeout 200 python bench.py --steps 5 --warmup 2 --workload h2_fluid_64000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c34_h64_s$S.json 2> gpurun_out/r02_c34_h64_s$S.err
done
python - <<'PY'
import json,glob,re
for f in sorted(glob.glob('gpurun_out/r02_c34_*.json'), key=lambda x:(x.split('_s')[0], int(re.findall(r'_s(\d+)',x)[0]))):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c34_')[1][:14].ljust(14), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()})
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
