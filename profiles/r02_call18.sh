#!/bin/bash
# round 2, GPU call 18 (one GPU): the no-allocation entry point alone in a fresh process, then the whole GPU suite (-x)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_c18_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c18_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c18_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c18_tests.log
tail -4 gpurun_out/r02_c18_noalloc_alone.log
tail -4 gpurun_out/r02_c18_tests.log
