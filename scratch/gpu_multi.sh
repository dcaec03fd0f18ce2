#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -15
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 200 --warmup 10 2> gpurun_out/bench2.err | tee gpurun_out/bench_n2.json
tail -5 gpurun_out/bench2.err
python bench.py --steps 200 --warmup 10 --no-cpu-baseline 2>/dev/null | tee gpurun_out/bench_n1.json
