#!/bin/bash
# SURVEY 8f f4 variants (Wolf electrostatics, rd_crystal image sums): bench lines, launch list and one --set full capture
# of crystal_lj_kernel; then the K7 pair-list captures (profiles/ncu_k7_list.sh).  One GPU, under gpurun.
#   bash profiles/ncu_f4.sh
mkdir -p gpurun_out
O=gpurun_out
python bench.py --workload co2_fluid_12k_rd_crystal2 > $O/r01_f4_co2_rd_crystal2.json 2> $O/r01_f4.err || exit 1
python bench.py --workload co2_fluid_12k_es_wolf --no-cpu-baseline > $O/r01_f4_co2_es_wolf.json 2>> $O/r01_f4.err
python bench.py --workload co2_fluid_12k_rd_crystal3 --no-cpu-baseline > $O/r01_f4_co2_rd_crystal3.json 2>> $O/r01_f4.err
python bench.py --workload mof_h2_63k_rd_crystal2 --no-cpu-baseline > $O/r01_f4_mof_rd_crystal2.json 2>> $O/r01_f4.err
W="--workload co2_fluid_12k_rd_crystal2 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $O/r01_f4_crystal_launches.csv \
    python bench.py --steps 1 --warmup 1 $W > $O/r01_f4_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:crystal_lj_kernel -s 1 -c 1 -o $O/r01_f4_crystal -f \
    python bench.py --steps 1 --warmup 1 $W > $O/r01_f4_ncu2.log 2>&1
ncu -i $O/r01_f4_crystal.ncu-rep --page details --csv > $O/r01_f4_crystal.details.csv 2>/dev/null
bash profiles/ncu_k7_list.sh
for f in $O/r01_f4_*.json; do echo $f; cut -c1-400 $f; done
