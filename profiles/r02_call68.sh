#!/bin/bash
# round 2, GPU call 68 (two GPUs): the sharded tests and the 2-GPU bench line with the one-image K4 and the new dispatch threshold
mkdir -p gpurun_out
nvidia-smi -L | head -4
timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_pins.py tests/test_gpu_shim.py -m gpu -q -x > gpurun_out/r02_c68_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c68_tests.log
tail -3 gpurun_out/r02_c68_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_c68_bench2.json 2> gpurun_out/r02_c68_bench2.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c68_bench2.json').read().strip().splitlines()[-1])
print('n_gpus', d['n_gpus'], 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'].get('value'), 'frac', d['roofline'].get('frac'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'))
PY
