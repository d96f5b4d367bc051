#!/bin/bash
# round 2, GPU call 59 (one GPU): drifted atoms relocated in the orders instead of a full sort: parity, large-move NVT A/B, whole suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "large_moves or small_edits" > gpurun_out/r02_c59_new_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c59_new_tests.log
tail -3 gpurun_out/r02_c59_new_tests.log
timeout 900 python profiles/big_moves_nvt.py mof_h2_21k 300 > gpurun_out/r02_c59_big_moves.txt 2>&1
timeout 900 python profiles/big_moves_nvt.py mof_h2_63k 60 >> gpurun_out/r02_c59_big_moves.txt 2>&1
cat gpurun_out/r02_c59_big_moves.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c59_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c59_tests.log
tail -3 gpurun_out/r02_c59_tests.log
