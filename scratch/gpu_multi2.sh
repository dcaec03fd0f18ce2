#!/bin/bash
# 2-GPU pass: multi-GPU parity tests, bench at N=2 (peer-memory and NCCL halo), N=1 config D (HR) line, adaptive test
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_rdy_driver.py -m gpu -x -q 2>&1 | tail -5
for halo in p2p nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 500 --warmup 20 --halo $halo 2> gpurun_out/bench_n2_$halo.err > gpurun_out/bench_n2_$halo.json; echo "n2 $halo exit $?"; cat gpurun_out/bench_n2_$halo.json; tail -3 gpurun_out/bench_n2_$halo.err
done
timeout 600 python bench.py --hr --steps 500 --warmup 20 --no-cpu-baseline 2> gpurun_out/bench_hr.err > gpurun_out/bench_n1_hr.json; echo "hr exit $?"; cat gpurun_out/bench_n1_hr.json
