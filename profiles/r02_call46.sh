#!/bin/bash
# round 2, GPU call 46 (four GPUs): multi-GPU tests and the 4-rank bench line of the final code
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_group.py tests/test_gpu_sharded.py -m gpu -q -x > gpurun_out/r02_c46_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c46_tests.log
tail -3 gpurun_out/r02_c46_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_c46_bench4.json 2> gpurun_out/r02_c46_bench4.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_c46_bench4.json').read().strip().splitlines()[-1])
print('n_gpus', d['n_gpus'], 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'].get('value'), 'parity', (d['config'].get('parity') or {}).get('max_rel_dev'), 'mc dropin', d['config'].get('mc_steps_per_s_dropin'))
PY
