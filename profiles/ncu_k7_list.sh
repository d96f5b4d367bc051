#!/bin/bash
# K7 pair-list form: launch list of one energy() + one --set full capture of the per-pass apply kernel and the fill kernel
#   bash profiles/ncu_k7_list.sh          (one GPU, under gpurun, after the plain run exited 0)
mkdir -p gpurun_out
W="--workload mof_h2_63k_ewald_full --no-cpu-baseline"
python bench.py --steps 3 --warmup 3 $W > gpurun_out/k7_list_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 200 --csv --log-file gpurun_out/r01_k7_list_launches.csv \
    python bench.py --steps 1 --warmup 3 $W > gpurun_out/k7_list_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"induced_list_apply_kernel|induced_list_kernel" -s 4 -c 3 \
    -o gpurun_out/r01_k7_list -f python bench.py --steps 1 --warmup 3 $W > gpurun_out/k7_list_ncu2.log 2>&1
ncu -i gpurun_out/r01_k7_list.ncu-rep --page details --csv > gpurun_out/r01_k7_list.details.csv 2>/dev/null
tail -1 gpurun_out/k7_list_plain.log
