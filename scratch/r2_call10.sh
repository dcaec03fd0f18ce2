#!/bin/bash
# round 2, GPU call 10 (1 GPU): K4 variants A/B (PM 1 vs 2), HR after the register fix, K3M first run + second-order A/B
cd /root/repo; mkdir -p gpurun_out
for v in 1 2; do
  RDYB_K4_PRIM=$v timeout 600 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab3_fo_pm$v.json 2>gpurun_out/r2_ab3_fo_pm$v.err; echo "fo pm=$v exit $?"
  RDYB_K4_PRIM=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_ab3_fo_pm${v}_s20.json 2>/dev/null; echo "fo pm=$v s20 exit $?"
done
timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab3_hr.json 2>/dev/null; echo "hr exit $?"
RDYB_K4_PRIM=2 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_regime.py -m gpu -q -x --timeout 600 > gpurun_out/r2_pytest_pm2.log 2>&1; echo "pytest pm2 exit $?"; tail -4 gpurun_out/r2_pytest_pm2.log
RDYB_TEST_UNVERIFIED=1 timeout 600 python -m pytest tests/test_gpu_muscl.py -m gpu -q --timeout 300 > gpurun_out/r2_pytest_muscl_unverified.log 2>&1; echo "pytest muscl exit $?"; tail -6 gpurun_out/r2_pytest_muscl_unverified.log
timeout 600 python bench.py --second-order --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_so_k2m.json 2>/dev/null; echo "so k2m exit $?"
RDYB_MUSCL_ASYNC=1 timeout 600 python bench.py --second-order --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_so_k3m.json 2>gpurun_out/r2_bench_so_k3m.err; echo "so k3m exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_ab3_*.json')+glob.glob('gpurun_out/r2_bench_so_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=d['checks']['parity'] or {}
        print(f.split('/')[-1], '%.4f ms  %.2f G  frac %.3f  parity max_rel %s ok %s'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], p.get('max_rel'), p.get('ok')), d['clocks']['reasons'], d['detail']['kernel_info'][:50])
    except Exception as e: print(f,'ERR',e)
PY
