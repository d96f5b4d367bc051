#!/bin/bash
# round 2, GPU call 79 (one GPU): tests/test_gpu_parity.py with the 3,072 threshold (what is left of the GPU budget)
mkdir -p gpurun_out
timeout 32 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r02_c79_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c79_tests.log
tail -3 gpurun_out/r02_c79_tests.log
