#!/bin/bash
# round 2, GPU call 75 (two GPUs): the sharded tests with the finer K4 items (small sharded cases now run with gridDim.z = 8)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q -x > gpurun_out/r02_c75_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c75_tests.log
tail -3 gpurun_out/r02_c75_tests.log
