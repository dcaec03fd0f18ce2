#!/bin/bash
# round 2, call 25 (last GPU seconds): K4M LEAN (36 KB stage, 3 CTAs/SM) - parity tests and one bench line
cd /root/repo; mkdir -p gpurun_out
export RDYB_K4M_LEAN=1
timeout 60 python -m pytest tests/test_gpu_muscl.py -q -x --timeout 50 -k "k4m" > gpurun_out/r2_k4m_lean_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r2_k4m_lean_pytest.log
timeout 60 python bench.py --second-order --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 1 > gpurun_out/r2_k4m_lean_so.json 2> gpurun_out/r2_k4m_lean_so.err; echo "lean exit $?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_k4m_lean_so.json').read().strip().splitlines()[-1]); print('lean %.4f ms %.2f G parity %s %s'%(d['ms_per_step'], d['value']/1e9, d['checks']['parity']['max_rel'], d['checks']['parity']['ok']))"
