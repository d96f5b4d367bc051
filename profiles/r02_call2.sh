#!/bin/bash
# round 2, GPU call 2 (two GPUs): the in-library sharded path -- group (one process) and one process per GPU -- and the rest of the suite
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_c2_topo.txt 2>&1
timeout 1200 python -m pytest tests/test_gpu_group.py tests/test_gpu_sharded.py tests/test_gpu_pins.py tests/test_gpu_shim.py tests/test_gpu_warm_start.py tests/test_gpu_incremental.py -m gpu -q > gpurun_out/r02_c2_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_c2_gputests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_c2_bench2.json 2> gpurun_out/r02_c2_bench2.err; echo "bench rc=$?" >> gpurun_out/r02_c2_bench2.err
MPMC_B200_STAGED=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c2_bench2_staged.json 2> gpurun_out/r02_c2_bench2_staged.err
tail -5 gpurun_out/r02_c2_gputests.log; cut -c1-400 gpurun_out/r02_c2_bench2.json; tail -3 gpurun_out/r02_c2_bench2.err
