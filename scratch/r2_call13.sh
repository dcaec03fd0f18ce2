#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bs2002.py -m gpu -q --timeout 300 > gpurun_out/r2_pytest_bs2002.log 2>&1; echo "pytest bs2002 exit $?"; tail -15 gpurun_out/r2_pytest_bs2002.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab5_fo_s20.json 2>/dev/null; echo "fo exit $?"
timeout 600 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_ab5_fo_s300.json 2>/dev/null; echo "fo300 exit $?"
timeout 600 python bench.py --hr --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab5_hr_s20.json 2>/dev/null; echo "hr exit $?"
timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_ab5_hr_s300.json 2>/dev/null; echo "hr300 exit $?"
timeout 600 python bench.py --second-order --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_ab5_so.json 2>/dev/null; echo "so exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_ab5*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=(d.get('checks') or {}).get('parity') or {}
        print(f.split('/')[-1], '%.4f ms  %.3f G  frac %.4f  parity max_rel %s ok %s'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], p.get('max_rel'), p.get('ok')), d['clocks']['reasons'])
    except Exception as e: print(f,'ERR',e)
PY
