#!/bin/bash
# round 2, GPU call 15 (two GPUs): partial-buffer cap (rounds), full suite on two GPUs, 63k bench sanity, large fluid
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_c15_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c15_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c15_bench.json 2> gpurun_out/r02_c15_bench.err
MPMC_B200_SYM_PART_MB=24 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c15_bench_cap24.json 2> gpurun_out/r02_c15_bench_cap24.err
timeout 400 python bench.py --steps 3 --warmup 2 --workload h2_fluid_256000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c15_bench_256k.json 2> gpurun_out/r02_c15_bench_256k.err
nvidia-smi --query-gpu=memory.used --format=csv > gpurun_out/r02_c15_mem.txt
tail -5 gpurun_out/r02_c15_tests.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c15_bench*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r02_c15_')[1][:22].ljust(22), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, (d['config'].get('parity') or {}).get('max_rel_dev'), d['roofline'].get('ms_per_launch'))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-500:])
PY
