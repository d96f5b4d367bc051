#!/bin/bash
# round 2, GPU call 10 (one GPU): adaptive K1 (walk inside the cutoff, compact across it), sqrt-free split rank metric
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02_c10_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c10_tests.log
for W in co2_fluid_12k h2_fluid_16000 h2_fluid_1000 mof_h2_21k; do
  timeout 200 python bench.py --steps 10 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c10_bench_$W.json 2> gpurun_out/r02_c10_bench_$W.err
done
timeout 200 python bench.py --steps 5 --warmup 3 --workload mof_h2_63k_gs_production --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c10_bench_gs.json 2> gpurun_out/r02_c10_bench_gs.err
tail -3 gpurun_out/r02_c10_tests.log
for f in gpurun_out/r02_c10_bench_*.json; do echo $f; cut -c1-200 $f; done
