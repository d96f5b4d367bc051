#!/bin/bash
# round 2, GPU call 76 (one GPU): one full ncu capture of the shipped K4 (one-image cluster pairs) on the bench workload
mkdir -p gpurun_out
timeout 60 python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c76_probe.log 2>&1 || { echo probe failed; tail -3 gpurun_out/r02_c76_probe.log; exit 1; }
timeout 80 ncu --set full --clock-control none --import-source on -k regex:dipole_sym_kernel -s 10 -c 1 -o gpurun_out/r02_k4_one_image -f python profiles/k4_probe.py mof_h2_63k > gpurun_out/r02_c76_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r02_k4_one_image.ncu-rep --page details --csv > gpurun_out/r02_k4_one_image_ncu_details.csv 2>/dev/null
rm -f gpurun_out/r02_k4_one_image.ncu-rep
grep -c . gpurun_out/r02_k4_one_image_ncu_details.csv
