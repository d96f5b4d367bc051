#!/bin/bash
# round 2, GPU call 16 (one GPU): K4 uniform-image shortcut, A/B against MPMC_B200_SYM_UNI=0; parity + pins
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pins.py -m gpu -q -x > gpurun_out/r02_c16_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c16_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c16_bench_uni.json 2> gpurun_out/r02_c16_bench_uni.err
MPMC_B200_SYM_UNI=0 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c16_bench_nouni.json 2> gpurun_out/r02_c16_bench_nouni.err
timeout 300 python bench.py --steps 10 --warmup 3 --workload mof_h2_21k --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c16_bench_21k.json 2> gpurun_out/r02_c16_bench_21k.err
timeout 300 python bench.py --steps 10 --warmup 3 --workload h2_fluid_16000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c16_bench_16k.json 2> gpurun_out/r02_c16_bench_16k.err
tail -5 gpurun_out/r02_c16_tests.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c16_bench*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c16_')[1][:22].ljust(22), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'), d['roofline'].get('ms_per_launch'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-500:])
PY
