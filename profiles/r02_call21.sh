#!/bin/bash
# round 2, GPU call 21 (one GPU): resident MC steps with the step replayed as a CUDA graph; tests, then A/B against plain launches
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_resident.py tests/test_gpu_shim.py -m gpu -q -x -k "resident or mc_steps" > gpurun_out/r02_c21_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c21_tests.log
tail -5 gpurun_out/r02_c21_tests.log
MPMC_B200_MC_GRAPH=1 timeout 900 python profiles/resident_steps_ab.py 0 16 64 > gpurun_out/r02_c21_resident_graph.txt 2>&1
MPMC_B200_MC_GRAPH=0 timeout 900 python profiles/resident_steps_ab.py 64 > gpurun_out/r02_c21_resident_nograph.txt 2>&1
echo "--- graph"; cat gpurun_out/r02_c21_resident_graph.txt; echo "--- plain launches"; cat gpurun_out/r02_c21_resident_nograph.txt
