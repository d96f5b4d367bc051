#!/bin/bash
# round 2, GPU call 50 (one GPU): carried orders with the new atoms placed next to their nearest survivor: set_atoms cost and the
# evaluation after it; uVT trajectories
mkdir -p gpurun_out
MPMC_B200_REPACK_TIMING=1 timeout 600 python profiles/set_atoms_cost.py > gpurun_out/r02_c50_set_atoms_carry.txt 2> gpurun_out/r02_c50_set_atoms_carry.err
cat gpurun_out/r02_c50_set_atoms_carry.txt; grep "n 21\|n 64" gpurun_out/r02_c50_set_atoms_carry.err | tail -2
timeout 900 python -m pytest tests/test_gpu_shim.py tests/test_gpu_parity.py -m gpu -q -x -k "uvt or trajectory or determinism or edge" > gpurun_out/r02_c50_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c50_tests.log
tail -3 gpurun_out/r02_c50_tests.log
