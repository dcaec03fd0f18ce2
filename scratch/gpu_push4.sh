#!/bin/bash
# 4 GPUs: 2x2 block partition parity (fused push default, up to 3 neighbours per rank), bench N=4 fused vs push kernel
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -k "four or peer_memory" 2>&1 | tail -4
for mode in fused one; do
  if [ $mode = one ]; then export RDYB_P2P_NO_FUSED_PUSH=1; else unset RDYB_P2P_NO_FUSED_PUSH; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 4 --steps 500 --warmup 20 2> gpurun_out/bench_n4_$mode.err > gpurun_out/bench_n4_$mode.json; echo "n4 $mode exit $?"
  python - <<PY
import json
j=json.loads(open("gpurun_out/bench_n4_$mode.json").read().strip().splitlines()[-1])
print("$mode", j["value"], j["ms_per_step"], j["gpu_launches"], j["clocks"])
PY
done
