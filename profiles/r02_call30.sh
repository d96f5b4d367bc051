#!/bin/bash
# round 2, GPU call 30 (eight GPUs): the 8-rank bench line of the final code (one process per GPU under torchrun), and the
# drop-in on a device group of 8
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_c30_bench8.json 2> gpurun_out/r02_c30_bench8.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c30_bench8.json').read().strip().splitlines()[-1])
print('n_gpus', d['n_gpus'], 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'].get('value'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'), 'mc dropin', d['config'].get('mc_steps_per_s_dropin'), 'frac', d['roofline'].get('frac'))
PY
