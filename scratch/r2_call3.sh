#!/bin/bash
# round 2, GPU call 3 (1 GPU): fp64 microbenchmark, launch list + ncu --set full of the fused-primitive first-order K4
cd /root/repo
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fp64_lat scratch/ubench/fp64_lat.cu && /tmp/fp64_lat > gpurun_out/r2_fp64_lat.txt 2>&1; cat gpurun_out/r2_fp64_lat.txt
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 --parity-steps 0"
$B > gpurun_out/plain_fo.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:swe_tiled_fixed -s 4 -c 1 -o gpurun_out/r2_prof_fixed_fo -f $B > gpurun_out/ncu_fo.log 2>&1; echo "ncu fo exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_bench_fixed.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 --parity-steps 0 > gpurun_out/ncu_l.log 2>&1; echo "ncu launches exit $?"
