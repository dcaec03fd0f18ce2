#!/bin/bash
# K6: parity on the triangle cases, bench, ncu capture
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "floodplain or rhs_apply" > gpurun_out/pytest_k6.log 2>&1
tail -3 gpurun_out/pytest_k6.log
timeout 300 python bench.py --kernel split --steps 300 --warmup 20 --no-cpu-baseline > gpurun_out/bench_k6.json 2> gpurun_out/bench_k6.err
tail -2 gpurun_out/bench_k6.err
python - <<'PY'
import json
for k in ("k6",):
    try:
        j=json.loads(open(f"gpurun_out/bench_{k}.json").read().strip().splitlines()[-1])
        print(k, j["value"], j["ms_per_step"], j["roofline"]["frac"], j["config"].get("kernel"))
    except Exception as e: print(k, "ERR", e)
PY
bash scratch/gpu_ncu.sh split 224 swe_tiled_split
