#!/bin/bash
# round 2, GPU call 80 (one GPU, the last seconds): the warm-start test, whose MOF case (3,438 polarizable atoms) changed dispatch
mkdir -p gpurun_out
timeout 14 python -m pytest tests/test_gpu_warm_start.py -m gpu -q -x > gpurun_out/r02_c80_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c80_tests.log
tail -3 gpurun_out/r02_c80_tests.log
