#!/bin/bash
# round 2, GPU call 55 (one GPU): subsets gathered on the device from the atom arrays (one upload per subset): set_atoms cost,
# uVT walk, whole GPU suite
mkdir -p gpurun_out
MPMC_B200_REPACK_TIMING=1 timeout 600 python profiles/set_atoms_cost.py > gpurun_out/r02_c55_set_atoms.txt 2> gpurun_out/r02_c55_set_atoms.err
cat gpurun_out/r02_c55_set_atoms.txt; grep "n 21\|n 64" gpurun_out/r02_c55_set_atoms.err | tail -2
timeout 600 python profiles/uvt_edit_walk.py mof_h2_21k 60 > gpurun_out/r02_c55_walk_21k.txt 2>&1; cat gpurun_out/r02_c55_walk_21k.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c55_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c55_tests.log
tail -3 gpurun_out/r02_c55_tests.log
