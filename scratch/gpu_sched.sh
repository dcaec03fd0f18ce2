#!/bin/bash
# one B200: A/B of the bank-aware edge order (device_layout.c tile_edge_schedule) on the bench kernel + parity subset
mkdir -p gpurun_out
for mode in sched nosched sched nosched; do
  if [ $mode = nosched ]; then export RDYB_NO_EDGE_SCHEDULE=1; else unset RDYB_NO_EDGE_SCHEDULE; fi
  timeout 600 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/bench_$mode.json 2> gpurun_out/bench_$mode.err; echo "$mode exit $?"
  python - <<PY
import json
j=json.loads(open("gpurun_out/bench_$mode.json").read().strip().splitlines()[-1])
print("$mode", j["value"], j["ms_per_step"], j["roofline"]["frac"], j["clocks"])
PY
done
unset RDYB_NO_EDGE_SCHEDULE
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_reference_decks.py -m gpu -q -x 2>&1 | tail -3
