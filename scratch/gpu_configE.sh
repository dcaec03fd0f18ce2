#!/bin/bash
# config E: ~100 M triangles over 8 GPUs (12.5 M per GPU: nx = 2500), peer-memory halo; plus the 2-rank second-order test
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 8 --nx 2500 --steps 300 --warmup 20 2> gpurun_out/bench_n8_configE.err > gpurun_out/bench_n8_configE.json; echo "configE exit $?"; cut -c1-600 gpurun_out/bench_n8_configE.json; tail -2 gpurun_out/bench_n8_configE.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29642 bench.py --gpus 8 --steps 300 --warmup 20 2> gpurun_out/bench_n8.err > gpurun_out/bench_n8.json; echo "n8 exit $?"; cut -c1-300 gpurun_out/bench_n8.json
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q 2>&1 | tail -3
