#!/bin/bash
# round 2, GPU call 49 (one GPU): spatial orders carried over a small edit of the atom set (uVT insert / remove): set_atoms cost A/B,
# then the whole GPU suite
mkdir -p gpurun_out
MPMC_B200_REUSE_ORDER=0 MPMC_B200_REPACK_TIMING=1 timeout 600 python profiles/set_atoms_cost.py > gpurun_out/r02_c49_set_atoms_full.txt 2> gpurun_out/r02_c49_set_atoms_full.err
MPMC_B200_REPACK_TIMING=1 timeout 600 python profiles/set_atoms_cost.py > gpurun_out/r02_c49_set_atoms_carry.txt 2> gpurun_out/r02_c49_set_atoms_carry.err
echo "--- full sort every time"; cat gpurun_out/r02_c49_set_atoms_full.txt; grep "n 21\|n 64" gpurun_out/r02_c49_set_atoms_full.err | tail -3
echo "--- orders carried over"; cat gpurun_out/r02_c49_set_atoms_carry.txt; grep "n 21\|n 64" gpurun_out/r02_c49_set_atoms_carry.err | tail -3
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c49_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c49_tests.log
tail -3 gpurun_out/r02_c49_tests.log
