#!/bin/bash
# round 2, GPU call 61 (one GPU): K4 with one image per cluster pair where the boxes prove it (no per-pair minimum image there):
# pins + parity of the K4-heavy tests, then A/B on the bench workload
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pins.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r02_c61_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c61_tests.log
tail -4 gpurun_out/r02_c61_tests.log
for u in 0 1 0 1; do
  MPMC_B200_SYM_UNISHIFT=$u timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c61_bench_u$u.json 2> gpurun_out/r02_c61_bench_u$u.err
  python - "$u" <<'PY'
import json,sys
u=sys.argv[1]
f='gpurun_out/r02_c61_bench_u%s.json'%u
try:
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print('unishift',u,'ms_per_step',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['roofline']['phase_ms'].items()},'frac',d['roofline'].get('frac'),'parity',(d['config'].get('parity') or {}).get('max_rel_dev'))
except Exception as e:
    print(f,'ERR',e,open(f.replace('.json','.err')).read()[-600:])
PY
done
for u in 0 1; do
  MPMC_B200_SYM_UNISHIFT=$u timeout 300 python bench.py --steps 10 --warmup 3 --workload mof_h2_21k --no-cpu-baseline --no-mc-dropin 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('mof_h2_21k unishift $u', round(d['ms_per_step'],3), (d['config'].get('parity') or {}).get('max_rel_dev'))"
done
