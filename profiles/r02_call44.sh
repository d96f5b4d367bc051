#!/bin/bash
# round 2, GPU call 44 (one GPU): the pin tests (incl. both K4 launch shapes) after the clean rebuild
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pins.py -m gpu -q -x > gpurun_out/r02_c44_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c44_tests.log
tail -3 gpurun_out/r02_c44_tests.log
