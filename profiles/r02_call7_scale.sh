#!/bin/bash
# round 2, GPU call 7 (eight GPUs of one box): scaling of one energy() at 1/2/4/8 GPUs on the SAME box (own arm only),
# the 8-rank sharded tests, a drop-in trajectory on a device group, the 8-chain ensemble workload
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_c7_topo.txt 2>&1
for N in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_c7_bench$N.json 2> gpurun_out/r02_c7_bench$N.err; echo "rc=$?" >> gpurun_out/r02_c7_bench$N.err
done
timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_c7_bench1.json 2> gpurun_out/r02_c7_bench1.err
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_sharded.py -m gpu -q > gpurun_out/r02_c7_multi_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_c7_multi_tests.log
timeout 300 python -m pytest tests/test_gpu_shim.py -m gpu -q -k "device_group" > gpurun_out/r02_c7_group_traj.log 2>&1
timeout 400 python bench.py --gpus 8 --steps 5 --warmup 3 --workload ensembles_h2_fluid_16000 --no-cpu-baseline > gpurun_out/r02_c7_bench_ens8.json 2> gpurun_out/r02_c7_bench_ens8.err
for N in 8 4 2 1; do cut -c1-260 gpurun_out/r02_c7_bench$N.json; done; tail -3 gpurun_out/r02_c7_multi_tests.log; tail -2 gpurun_out/r02_c7_group_traj.log; cut -c1-300 gpurun_out/r02_c7_bench_ens8.json
