#!/bin/bash
# round 2, GPU call 11 (four GPUs): drop-in energy(system_t*) on a device group with / without per-rank host threads; 4 chains
mkdir -p gpurun_out
python - > gpurun_out/r02_c11_dropin_group_timing.txt 2>&1 <<'PY'
import os, subprocess, sys, tempfile
sys.path.insert(0, os.getcwd())
import bench
from mpmc_b200 import systems
s = bench.build_workload("mof_h2_63k")
with tempfile.TemporaryDirectory() as td:
    inp = systems.write_reference_case(s, td)
    for env in ({"MPMC_B200_DEVICES": "0-3"}, {"MPMC_B200_DEVICES": "0-3", "MPMC_B200_GROUP_THREADS": "0"}, {"MPMC_B200_DEVICES": "0,1"}, {}):
        r = subprocess.run([bench.SHIM_DRIVER, "input", "30", "5", "q"], cwd=td, capture_output=True, text=True,
                           env=dict(os.environ, MPMC_B200_SHIM_TIMING="1", **env))
        print(env, [l for l in r.stdout.splitlines() if l.startswith("timing")], r.stderr[-400:])
PY
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29704 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_c11_bench4.json 2> gpurun_out/r02_c11_bench4.err
timeout 400 python bench.py --gpus 4 --steps 5 --warmup 3 --workload ensembles_h2_fluid_16000 --no-cpu-baseline > gpurun_out/r02_c11_bench_ens4.json 2> gpurun_out/r02_c11_bench_ens4.err
timeout 600 python -m pytest tests/test_gpu_group.py tests/test_gpu_sharded.py -m gpu -q > gpurun_out/r02_c11_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02_c11_tests.log
cat gpurun_out/r02_c11_dropin_group_timing.txt; cut -c1-300 gpurun_out/r02_c11_bench4.json; cut -c1-200 gpurun_out/r02_c11_bench_ens4.json; tail -3 gpurun_out/r02_c11_tests.log
