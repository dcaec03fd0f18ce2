#!/bin/bash
cd /root/repo
for t in 224 160; do
timeout 300 python bench.py --kernel fixed --tile $t --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_t$t.json 2> gpurun_out/bench_t$t.err
python - <<PY
import json
j=json.loads(open("gpurun_out/bench_t$t.json").read().strip().splitlines()[-1])
print($t, j["value"], j["ms_per_step"], j["roofline"]["frac"], j["config"]["kernel_info"][:120])
PY
done
