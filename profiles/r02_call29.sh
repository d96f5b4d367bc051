#!/bin/bash
# round 2, GPU call 29 (one GPU): resident MC steps re-measured with the corrected interval (1st .. last call, so the first batch
# is inside it), ring of 512 MB so that the large systems get full batches
mkdir -p gpurun_out
timeout 1500 python profiles/resident_steps_ab.py 0 16 64 > gpurun_out/r02_c29_resident_ab.txt 2>&1
cat gpurun_out/r02_c29_resident_ab.txt
