#!/bin/bash
# round 2, GPU call 25 (one GPU): K1 arithmetic (rsqrt / per-atom sqrt(eps), 1/m / erfc_fast): the whole GPU suite
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c25_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c25_tests.log
tail -5 gpurun_out/r02_c25_tests.log
