#!/bin/bash
# round 2, GPU call 71 (one GPU): an evaluation right after an edit of the atom set against the steady one (7k-atom MOF, 4 passes)
mkdir -p gpurun_out
MPMC_B200_REPACK_TIMING=1 timeout 300 python profiles/energy_after_edit.py mof_h2_7k 4 > gpurun_out/r02_c71_after_edit.txt 2> gpurun_out/r02_c71_after_edit.err
cat gpurun_out/r02_c71_after_edit.txt; grep -c carried gpurun_out/r02_c71_after_edit.err; tail -3 gpurun_out/r02_c71_after_edit.err
