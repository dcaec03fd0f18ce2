#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
( timeout 300 $T scratch/debug_hr_parity.py --nx 1200 --steps 3
  timeout 400 $T scratch/debug_hr_parity.py --nx 3536 --ny 1200 --steps 3
  timeout 400 $T scratch/debug_hr_parity.py --nx 3536 --ny 1200 --steps 2 --kernel 0 --halo nccl
  timeout 400 $T scratch/debug_hr_parity.py --nx 3536 --ny 1200 --steps 2 --no-dam
) > gpurun_out/r2_debug_hr.log 2>&1
grep -v "^W\|^\*\*\*\|Setting OMP\|warn" gpurun_out/r2_debug_hr.log | tail -70
# A/B of the split barrier (new default) on one GPU meanwhile
for v in 0 1; do
  RDYB_K4_PRIM_PASS=$v CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab2_fo_prim$v.json 2>/dev/null; echo "fo prim_pass=$v exit $?"
  RDYB_K4_PRIM_PASS=$v CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_ab2_hr_prim$v.json 2>/dev/null; echo "hr prim_pass=$v exit $?"
done
CUDA_VISIBLE_DEVICES=0 timeout 600 python -m pytest tests/test_rdy_driver.py tests/test_gpu_parity.py -m gpu -q -x --timeout 600 -k "device_side or floodplain or dam_break" 2>&1 | tail -5
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_ab2_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=d['checks']['parity']
        print(f, '%.4f ms  %.2f G  frac %.3f  parity max_rel %.2e normwise %.2e ok %s'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], p['max_rel'], p['normwise'], p['ok']), d['clocks'])
    except Exception as e: print(f,'ERR',e)
PY
