#!/bin/bash
# round 2, GPU call 60 (one GPU): the largest sweep point of BASELINE config 5 (256k atoms) and the default bench line of the final code
mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --warmup 2 --workload h2_fluid_256000 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c60_bench_256k.json 2> gpurun_out/r02_c60_bench_256k.err
timeout 600 python bench.py > gpurun_out/r02_c60_bench.json 2> gpurun_out/r02_c60_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/r02_c60_bench_256k.json','gpurun_out/r02_c60_bench.json'):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'ms_per_step', round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['phase_ms'].items()}, 'value', round(d['value'],2), 'e2e', d['e2e'].get('value'), 'frac', d['roofline'].get('frac'), (d['config'].get('parity') or {}).get('max_rel_dev'))
    except Exception as e:
        print(f,'ERR',e, open(f.replace('.json','.err')).read()[-400:])
PY
