#!/bin/bash
# round 2, call 20: K4M (fixed-shape second-order tile kernel with in-kernel gradients) - parity tests, then A/B bench lines
cd /root/repo; mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_muscl.py tests/test_gpu_bench_regime.py -q -x --timeout 300 -k "k4m or second_order_one_million or async_tiled" > gpurun_out/r2_k4m_pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/r2_k4m_pytest.log
B="timeout 200 python bench.py --second-order --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 1"
$B > gpurun_out/r2_k4m_so.json 2> gpurun_out/r2_k4m_so.err; echo "k4m exit $?"
RDYB_K4M_CONSUMERS=256 $B --parity-steps 0 > gpurun_out/r2_k4m_so_c256.json 2> gpurun_out/r2_k4m_so_c256.err; echo "c256 exit $?"
RDYB_K4M_CONSUMERS=224 $B --parity-steps 0 > gpurun_out/r2_k4m_so_c224.json 2> gpurun_out/r2_k4m_so_c224.err; echo "c224 exit $?"
RDYB_K4M_GRAD_ARRAY=1 $B --parity-steps 0 > gpurun_out/r2_k4m_so_gradarray.json 2> gpurun_out/r2_k4m_so_gradarray.err; echo "gradarray exit $?"
RDYB_MUSCL_K3M=1 $B --parity-steps 0 > gpurun_out/r2_k4m_so_k3m.json 2> gpurun_out/r2_k4m_so_k3m.err; echo "k3m exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_k4m_so*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=(d.get('checks') or {}).get('parity') or {}
        print(f.split('/')[-1], '%.4f ms %.2f G frac %.3f'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'launches', d['gpu_launches'], 'parity', p.get('max_rel'), p.get('ok'), d['detail']['kernel_info'][:70], d['clocks']['reasons'])
    except Exception as e: print(f, 'unreadable', e); print(open(f.replace('.json','.err')).read()[-600:])
PY
