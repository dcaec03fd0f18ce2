#!/bin/bash
# 8-GPU pass: 4-rank parity test, bench at N=8 and N=4 (weak scaling, peer-memory halo), N=8 NCCL for comparison
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -k four 2>&1 | tail -3
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 500 --warmup 20 2> gpurun_out/bench_n$n.err > gpurun_out/bench_n$n.json; echo "n$n exit $?"; cat gpurun_out/bench_n$n.json | cut -c1-400; tail -2 gpurun_out/bench_n$n.err
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29627 bench.py --gpus 8 --steps 300 --warmup 20 --halo nccl 2> gpurun_out/bench_n8_nccl.err > gpurun_out/bench_n8_nccl.json; echo "n8 nccl exit $?"; cat gpurun_out/bench_n8_nccl.json | cut -c1-300
