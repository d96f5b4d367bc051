#!/bin/bash
# round 2, GPU call 27 (one GPU): rint of the minimum image in K1 / K3 / K7 -- two DADDs (magic number) against FRND.F64 on the
# conversion pipe.  A/B by rebuilding the library on the box; K1 and K3 timed without their overlap (MPMC_B200_OVERLAP=0).
mkdir -p gpurun_out
run() {
  for W in co2_fluid_12k h2_fluid_16000 h2_fluid_64000 mof_h2_63k mof_h2_63k_ewald_full; do
    MPMC_B200_OVERLAP=0 timeout 300 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c27_$1_$W.json 2> gpurun_out/r02_c27_$1_$W.err
  done
}
run magic
touch mpmc_b200/csrc/common.cuh
make -s -j16 -C mpmc_b200/csrc EXTRA=-DMPMC_RINT_FRND > gpurun_out/r02_c27_build.log 2>&1; echo "build rc=$?"
run frnd
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pins.py -m gpu -q -x > gpurun_out/r02_c27_tests_frnd.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c27_tests_frnd.log
tail -3 gpurun_out/r02_c27_tests_frnd.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c27_*_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c27_')[1][:34].ljust(34), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
