#!/bin/bash
# usage: gpu_sweep.sh "<kernel:tile> ..." [extra bench args]   (no pytest)
mkdir -p gpurun_out
bash scratch/sweep.sh "$1" ${@:2} 2>&1 | tee gpurun_out/sweep.log
