#!/bin/bash
# round 2, GPU call 58 (one GPU): NVT through the drop-in with large displacement moves (drift re-sorts)
mkdir -p gpurun_out
timeout 900 python profiles/big_moves_nvt.py mof_h2_21k 300 > gpurun_out/r02_c58_big_moves.txt 2>&1
timeout 900 python profiles/big_moves_nvt.py mof_h2_63k 60 >> gpurun_out/r02_c58_big_moves.txt 2>&1
cat gpurun_out/r02_c58_big_moves.txt
