#!/bin/bash
# round 2, GPU call 78 (one GPU, last of the round): symmetric K4 from 3,072 padded polarizable atoms: the mid-size oracle cases
mkdir -p gpurun_out
timeout 50 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "test_cuda_matches_oracle_midsize" > gpurun_out/r02_c78_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c78_tests.log
tail -3 gpurun_out/r02_c78_tests.log
