#!/bin/bash
cd /root/repo
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_mms.py tests/test_reference_decks.py tests/test_gpu_muscl.py -m gpu -q > gpurun_out/pytest_roe.log 2>&1
tail -5 gpurun_out/pytest_roe.log
timeout 300 python bench.py --kernel fixed --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_roe.json 2> gpurun_out/bench_roe.err
python - <<PY
import json
j=json.loads(open("gpurun_out/bench_roe.json").read().strip().splitlines()[-1])
print("K4", j["value"], j["ms_per_step"], j["roofline"]["frac"])
PY
