#!/bin/bash
cd /root/repo
RDYB_DEBUG_COPYONLY=1 timeout 300 python bench.py --kernel fixed --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_copyonly.json 2> gpurun_out/bench_copyonly.err
timeout 300 python bench.py --kernel fixed --no-courant --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_nocourant.json 2> gpurun_out/bench_nocourant.err
RDYB_CTAS_PER_SM=2 timeout 300 python bench.py --kernel fixed --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_2cta.json 2> gpurun_out/bench_2cta.err
RDYB_CTAS_PER_SM=1 timeout 300 python bench.py --kernel fixed --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_1cta.json 2> gpurun_out/bench_1cta.err
python - <<PY
import json
for k in ("copyonly","nocourant","2cta","1cta"):
    try:
        j=json.loads(open(f"gpurun_out/bench_{k}.json").read().strip().splitlines()[-1])
        print(k, j["value"], j["ms_per_step"], j["roofline"]["frac"])
    except Exception as e: print(k, "ERR", e)
PY
