#!/bin/bash
# one B200: full GPU test suite, smoke, default bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/ -m gpu -q > gpurun_out/pytest_gpu_full.log 2>&1; tail -4 gpurun_out/pytest_gpu_full.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench.json
