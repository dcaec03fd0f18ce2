#!/bin/bash
# 2 GPUs: multi-GPU parity tests with the fused in-kernel push, bench N=2 fused push vs one-launch with a push kernel
mkdir -p gpurun_out
RDYB_P2P_FUSED_PUSH=1 timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q 2>&1 | tail -6
for mode in fused one; do
  if [ $mode = fused ]; then export RDYB_P2P_FUSED_PUSH=1; else unset RDYB_P2P_FUSED_PUSH; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29653 bench.py --gpus 2 --steps 500 --warmup 20 2> gpurun_out/bench_n2_$mode.err > gpurun_out/bench_n2_$mode.json; echo "n2 $mode exit $?"
  python - <<PY
import json
j=json.loads(open("gpurun_out/bench_n2_$mode.json").read().strip().splitlines()[-1])
print("$mode", j["value"], j["ms_per_step"], j["gpu_launches"], j["config"]["kernel_info"][:110])
PY
done
