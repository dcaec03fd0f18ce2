#!/bin/bash
cd /root/repo
for e in 60 0.5; do
timeout 300 python bench.py --kernel fixed --eta0 $e --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_eta$e.json 2> gpurun_out/bench_eta$e.err
python - <<PY
import json
j=json.loads(open("gpurun_out/bench_eta$e.json").read().strip().splitlines()[-1])
print("eta0", $e, j["value"], j["ms_per_step"], j["roofline"]["frac"], j["checks"])
PY
done
