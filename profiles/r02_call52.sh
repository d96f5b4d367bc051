#!/bin/bash
# round 2, GPU call 52 (two GPUs): whole GPU suite (incl. the device-group and one-process-per-GPU tests) with the carried orders
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c52_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c52_tests.log
tail -3 gpurun_out/r02_c52_tests.log
