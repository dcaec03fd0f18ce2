#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q --timeout 300 -k "rk4 or second_order or nccl" > gpurun_out/r2_pytest_gpu_multi2_rk4.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/r2_pytest_gpu_multi2_rk4.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
timeout 600 $T bench.py --gpus 2 --rk4 --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 1 --parity-steps 0 > gpurun_out/r2_bench_n2_rk4_p2p.json 2>/dev/null; echo "rk4 p2p exit $?"
RDYB_RK4_NCCL=1 timeout 600 $T bench.py --gpus 2 --rk4 --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 1 --parity-steps 0 > gpurun_out/r2_bench_n2_rk4_nccl.json 2>/dev/null; echo "rk4 nccl exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench_n2_rk4*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], '%.4f ms  %.3f G  frac %.4f'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), d['detail']['per_rank_ms_per_step'])
    except Exception as e: print(f,'ERR',e)
PY
