#!/bin/bash
# round 2, call 22: K4M with the packed ring-1 record (A/B), parity tests
cd /root/repo; mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_muscl.py tests/test_gpu_bench_regime.py -q --timeout 300 -k "k4m or second_order_one_million" > gpurun_out/r2_k4m_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r2_k4m_pytest.log
B="timeout 200 python bench.py --second-order --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 1"
$B > gpurun_out/r2_k4m_so.json 2> gpurun_out/r2_k4m_so.err; echo "k4m exit $?"
RDYB_K4M_NO_HALO_REC=1 $B --parity-steps 0 > gpurun_out/r2_k4m_so_norec.json 2> gpurun_out/r2_k4m_so_norec.err; echo "k4m norec exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_k4m_so.json')+glob.glob('gpurun_out/r2_k4m_so_norec.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=(d.get('checks') or {}).get('parity') or {}
        print(f.split('/')[-1], '%.4f ms %.2f G frac %.3f'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'launches', d['gpu_launches'], 'parity', p.get('max_rel'), p.get('ok'), d['detail']['kernel_info'][:70], d['clocks']['reasons'])
    except Exception as e: print(f, 'unreadable', e); print(open(f.replace('.json','.err')).read()[-600:])
PY
