#!/bin/bash
# round 2, GPU call 20 (one GPU): resident MC steps, MC steps/s A/B measured by the shim's own clock
mkdir -p gpurun_out
timeout 1500 python profiles/resident_steps_ab.py 0 8 32 64 > gpurun_out/r02_c20_resident_ab.txt 2>&1
cat gpurun_out/r02_c20_resident_ab.txt
