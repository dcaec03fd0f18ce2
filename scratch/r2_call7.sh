#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
( timeout 600 $T scratch/debug_hr_parity.py --nx 3536 --steps 5 ) > gpurun_out/r2_debug_hr2.log 2>&1 &
CUDA_VISIBLE_DEVICES=0 true
wait
grep -v "^W\|^\*\*\*\|Setting OMP\|warn" gpurun_out/r2_debug_hr2.log | tail -40
CUDA_VISIBLE_DEVICES=0 timeout 900 python bench.py --hr --nx 3536 --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 1 > gpurun_out/r2_bench_hr_25M_n1.json 2> gpurun_out/r2_bench_hr_25M_n1.err; echo "hr 25M n1 exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_hr_25M_n1.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], json.dumps(d['checks']['parity']))
PY
