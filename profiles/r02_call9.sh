#!/bin/bash
# round 2, GPU call 9 (one GPU): compacted K1, split palmo rows, launch list of the GS production evaluation
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c9_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c9_tests.log
for W in co2_fluid_12k h2_fluid_16000 mof_h2_63k; do
  timeout 200 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c9_bench_$W.json 2> gpurun_out/r02_c9_bench_$W.err
  MPMC_B200_K1=walk timeout 200 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c9_bench_${W}_walk.json 2> gpurun_out/r02_c9_bench_${W}_walk.err
done
timeout 200 python bench.py --steps 3 --warmup 2 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c9_bench_gs.json 2> gpurun_out/r02_c9_bench_gs.err
timeout 200 python profiles/k4_probe.py mof_h2_63k_gs_production > gpurun_out/r02_c9_gs_probe.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_c9_launches_gs.csv python profiles/k4_probe.py mof_h2_63k_gs_production > gpurun_out/r02_c9_ncu.log 2>&1
tail -3 gpurun_out/r02_c9_tests.log
for f in gpurun_out/r02_c9_bench_*.json; do echo $f; cut -c1-220 $f; done
