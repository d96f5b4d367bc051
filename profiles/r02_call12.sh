#!/bin/bash
# round 2, GPU call 12 (one GPU): fused small-system solver -- full suite + N sweep with and without it
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02_c12_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c12_tests.log
for W in h2_fluid_1000 h2_fluid_2000 h2_fluid_4000 h2_fluid_8000 mof_h2_7k; do
  timeout 200 python bench.py --steps 20 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c12_bench_$W.json 2> gpurun_out/r02_c12_bench_$W.err
  MPMC_B200_FUSED_MAX=0 timeout 200 python bench.py --steps 20 --warmup 3 --workload $W --no-cpu-baseline --no-mc-dropin > gpurun_out/r02_c12_bench_${W}_nofused.json 2> gpurun_out/r02_c12_bench_${W}_nofused.err
done
python - > gpurun_out/r02_c12_small_dropin.txt 2>&1 <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import bench
from tests import golden_util as gu
s, _ = gu.load("ref_socmof")
first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
s.meta["atom_mass"] = np.where(first, s.mol_mass, 0.0)
for env in ({}, {"MPMC_B200_FUSED_MAX": "0"}):
    os.environ.update(env)
    print("socMOF+BSSP 1,236 atoms NVT", env, "MC steps/s through mpmc_b200 <input>:", bench.dropin_mc_steps_per_s(s, 1, 100, 4100))
s, _ = gu.load("ref_small")
s.meta["atom_mass"] = np.where(np.r_[True, s.mol_id[1:] != s.mol_id[:-1]], s.mol_mass, 0.0)
os.environ.pop("MPMC_B200_FUSED_MAX", None)
print("cuda_pol.small 10 atoms NVT MC steps/s:", bench.dropin_mc_steps_per_s(s, 1, 200, 8200))
PY
tail -4 gpurun_out/r02_c12_tests.log; for f in gpurun_out/r02_c12_bench_*.json; do echo $f; cut -c60-200 $f; done; cat gpurun_out/r02_c12_small_dropin.txt
