#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest.log
timeout 600 python bench.py --no-cpu-baseline --steps 300 > gpurun_out/bench_pipe.json 2> gpurun_out/bench_pipe.err; echo "bench exit $?"; python -c "
import json; j=json.loads(open('gpurun_out/bench_pipe.json').read().strip().splitlines()[-1]); print(j['value'], j['roofline']['frac'], j['e2e'])"
RDYB_HOST_PIPELINE=0 timeout 600 python bench.py --no-cpu-baseline --steps 100 > gpurun_out/bench_nopipe.json 2>/dev/null; python -c "
import json; j=json.loads(open('gpurun_out/bench_nopipe.json').read().strip().splitlines()[-1]); print('nopipe', j['e2e'])"
for n in 4 8 32; do RDYB_HOST_PIPELINE=$n timeout 600 python bench.py --no-cpu-baseline --steps 100 > gpurun_out/bench_pipe$n.json 2>/dev/null; python -c "
import json; j=json.loads(open('gpurun_out/bench_pipe$n.json').read().strip().splitlines()[-1]); print('chunks $n', j['e2e']['ms_per_step'])"; done
