#!/bin/bash
# round 2, GPU call 45 (one GPU): K4 resident CTAs per SM -- __launch_bounds__(128, 4) (128 registers) against 5 (96 registers, 40 B
# of spills) and 6; the library rebuilt on the box for each
mkdir -p gpurun_out
run() {
  for W in mof_h2_63k h2_fluid_16000; do
    timeout 300 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c45_$1_$W.json 2> gpurun_out/r02_c45_$1_$W.err
  done
}
run minb4
for M in 5 6; do
  touch mpmc_b200/csrc/dipole_sym_kernel.cu
  make -s -j16 -C mpmc_b200/csrc EXTRA=-DSYM_MINB=$M > gpurun_out/r02_c45_build$M.log 2>&1; echo "build $M rc=$?"
  run minb$M
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c45_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c45_')[1][:28].ljust(28), round(d['ms_per_step'],3), 'K4/launch', round(d['roofline'].get('ms_per_launch') or 0,4), (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
