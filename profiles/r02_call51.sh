#!/bin/bash
# round 2, GPU call 51 (one GPU): uVT-like walk of random molecule removals / insertions: carried orders against a full sort every time
mkdir -p gpurun_out
for W in mof_h2_21k mof_h2_63k; do
  timeout 600 python profiles/uvt_edit_walk.py $W 120 > gpurun_out/r02_c51_walk_${W}_carry.txt 2>&1
  MPMC_B200_REUSE_ORDER=0 timeout 600 python profiles/uvt_edit_walk.py $W 60 > gpurun_out/r02_c51_walk_${W}_full.txt 2>&1
  echo "--- $W carried"; cat gpurun_out/r02_c51_walk_${W}_carry.txt; echo "--- $W full sort"; cat gpurun_out/r02_c51_walk_${W}_full.txt
done
