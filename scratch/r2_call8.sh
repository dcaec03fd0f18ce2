#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
( timeout 600 $T scratch/debug_hr_parity.py --nx 3536 --steps 5 --multi
  timeout 600 $T scratch/debug_hr_parity.py --nx 3536 --steps 5 --multi --fo
  timeout 600 $T scratch/debug_hr_parity.py --nx 1200 --steps 5 --multi ) > gpurun_out/r2_debug_hr3.log 2>&1
grep -v "^W\|^\*\*\*\|Setting OMP\|warn" gpurun_out/r2_debug_hr3.log | tail -40
