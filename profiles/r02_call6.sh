#!/bin/bash
# round 2, GPU call 6 (one GPU): full suite with the PCN-61 uVT averages test; ncu --set full of the persistent GS sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_c6_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_c6_gputests.log
timeout 300 python profiles/k4_shard_probe.py 8 4 --one > gpurun_out/r02_c6_shard_probe.txt 2>&1
timeout 300 python profiles/k4_probe.py mof_h2_63k_gs_production > gpurun_out/r02_c6_gs_probe.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gs_sweep_persistent -s 5 -c 1 -o gpurun_out/r02_gs_sweep -f python profiles/k4_probe.py mof_h2_63k_gs_production > gpurun_out/r02_c6_ncu.log 2>&1
echo "ncu rc=$?" >> gpurun_out/r02_c6_ncu.log
tail -5 gpurun_out/r02_c6_gputests.log; cat gpurun_out/r02_c6_shard_probe.txt; tail -3 gpurun_out/r02_c6_ncu.log; ls -la gpurun_out/*.ncu-rep
