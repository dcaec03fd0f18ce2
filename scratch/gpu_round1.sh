#!/bin/bash
# first GPU pass of the session: parity tests, bench (both arms), ncu launch list + full capture of the top kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
echo "ncu launches exit $?"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:swe_tiled_async -s 4 -c 2 -o gpurun_out/prof_async $B > gpurun_out/ncu2.log 2>&1
echo "ncu full exit $?"
