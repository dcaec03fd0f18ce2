#!/bin/bash
# round 2, GPU call 1 (1 GPU): full GPU suite with the capped-grid variants + 1M-cell tests, smoke, bench with checks.parity, HR bench, ncu
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -m gpu -q -x --timeout 600 > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/r2_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/r2_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench exit $?"; cut -c1-400 gpurun_out/r2_bench_n1.json; tail -3 gpurun_out/r2_bench_n1.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref exit $?"; cut -c1-200 gpurun_out/r2_bench_ref.json
timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_hr.json 2> gpurun_out/r2_bench_hr.err; echo "hr exit $?"; cut -c1-300 gpurun_out/r2_bench_n1_hr.json
timeout 600 python bench.py --eta0 60 --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_wet.json 2> gpurun_out/r2_bench_wet.err; echo "wet exit $?"
timeout 600 python bench.py --rk4 --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_rk4.json 2> gpurun_out/r2_bench_rk4.err; echo "rk4 exit $?"
B="python bench.py --hr --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 --parity-steps 0"
$B > gpurun_out/plain_hr.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:swe_tiled_fixed -s 4 -c 1 -o gpurun_out/r2_prof_fixed_hr -f $B > gpurun_out/ncu_hr.log 2>&1; echo "ncu hr exit $?"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
