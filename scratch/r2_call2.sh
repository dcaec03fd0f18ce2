#!/bin/bash
# round 2, GPU call 2 (1 GPU): full GPU suite (fused-primitive K4 is the default now), A/B of the two K4 variants, first order and HR
cd /root/repo
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -m gpu -q --timeout 600 > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/r2_pytest_gpu.log
for v in 0 1; do
  RDYB_K4_PRIM_PASS=$v timeout 600 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab_fo_prim$v.json 2> gpurun_out/r2_ab_fo_prim$v.err; echo "fo prim_pass=$v exit $?"
  RDYB_K4_PRIM_PASS=$v timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab_hr_prim$v.json 2> gpurun_out/r2_ab_hr_prim$v.err; echo "hr prim_pass=$v exit $?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_ab_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=d['checks']['parity']
        print(f, '%.4f ms  %.2f G  frac %.3f  parity max_rel %.2e normwise %.2e ok %s'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], p['max_rel'], p['normwise'], p['ok']), d['clocks'])
    except Exception as e: print(f,'ERR',e)
PY
