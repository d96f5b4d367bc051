#!/bin/bash
# round 2, GPU call 17 (one GPU): the no-allocation entry point (mpmc_b200_recip_energy_static) alone in a fresh process,
# then the whole GPU suite the way the driver runs it (-x), then the default bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_c17_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c17_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c17_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c17_tests.log
timeout 600 python bench.py > gpurun_out/r02_c17_bench.json 2> gpurun_out/r02_c17_bench.err
tail -4 gpurun_out/r02_c17_noalloc_alone.log
tail -4 gpurun_out/r02_c17_tests.log
tail -c 1500 gpurun_out/r02_c17_bench.json
