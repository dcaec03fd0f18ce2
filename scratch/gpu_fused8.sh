#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -k four 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29671 bench.py --gpus 8 --steps 500 --warmup 20 2> gpurun_out/bench_n8.err > gpurun_out/bench_n8.json; echo "n8 exit $?"
python - <<PY
import json
j=json.loads(open("gpurun_out/bench_n8.json").read().strip().splitlines()[-1])
print("n8", j["value"], j["ms_per_step"], j["gpu_launches"], j["roofline"]["frac"], j["clocks"])
PY
