#!/bin/bash
# round-end pass on one B200: full GPU test suite, smoke, bench (both arms), launch list, ncu full capture of the bench kernel
cd /root/repo
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/ -m gpu -q > gpurun_out/pytest_gpu_full.log 2>&1; tail -4 gpurun_out/pytest_gpu_full.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench.json
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; cut -c1-300 gpurun_out/bench_ref.json
timeout 600 python bench.py --second-order --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 > gpurun_out/bench_muscl.json 2> gpurun_out/bench_muscl.err; echo "muscl exit $?"
timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/bench_n1_hr.json 2> gpurun_out/bench_hr.err; echo "hr exit $?"
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1; echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:swe_tiled_fixed -s 4 -c 1 -o gpurun_out/prof_fixed -f $B > gpurun_out/ncu2.log 2>&1; echo "ncu full exit $?"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
