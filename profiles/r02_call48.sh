#!/bin/bash
# round 2, GPU call 48 (one GPU): cost of set_atoms (uVT insertion / removal) beside energy()
mkdir -p gpurun_out
timeout 600 python profiles/set_atoms_cost.py > gpurun_out/r02_c48_set_atoms.txt 2>&1; cat gpurun_out/r02_c48_set_atoms.txt
