#!/bin/bash
# usage: gpu_ncu.sh <kernel> <tile> <regex> [extra bench args]
mkdir -p gpurun_out
B="python bench.py --kernel $1 --tile $2 --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 ${@:4}"
$B > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$3 -s 4 -c 1 -o gpurun_out/prof_$1_$2 -f $B > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -3 gpurun_out/ncu_full.log
