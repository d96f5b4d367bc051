#!/bin/bash
# A/B of the multi-GPU switches on one 8-GPU box (results: profiles/r01_multi_gpu.md).
#   bash profiles/ab_multi_gpu.sh            # under: gpurun --gpus 8
run() { # n overlap peer hsplit
  MPMC_B200_OVERLAP=$2 MPMC_B200_PEER=$3 MPMC_B200_SYM_HSPLIT=$4 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 295$1$2 bench.py --gpus $1 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -n1 > gpurun_out/ab_tmp.log
  python - <<PY
import json
d=json.loads(open("gpurun_out/ab_tmp.log").read())
r=d["roofline"]
print("n=$1 overlap=$2 peer=$3 hsplit=$4 ms", round(d["ms_per_step"],3), "e2e_steps/s", round(d["e2e"]["mc_steps_per_s"],2), "k4", round(r["ms_per_launch"],4), {k:round(v,3) for k,v in r["phase_ms"].items()}, flush=True)
PY
}
mkdir -p gpurun_out
for cfg in "$@"; do run $cfg; done
