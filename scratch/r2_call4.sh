#!/bin/bash
# round 2, GPU call 4 (2 GPUs): multi-GPU parity tests (fused push, caps, mixed paths, HR, second order), adapter tests, N=2 bench lines, config D at 2 GPUs
cd /root/repo
mkdir -p gpurun_out
RDYB_TEST_UNVERIFIED=1 timeout 1500 python -m pytest tests/test_multi_gpu.py tests/test_gpu_rdycore_adapter.py -m gpu -q --timeout 600 > gpurun_out/r2_pytest_gpu_multi2.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/r2_pytest_gpu_multi2.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
timeout 900 $T bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2_s20.json 2> gpurun_out/r2_bench_n2_s20.err; echo "n2 s20 exit $?"
timeout 900 $T bench.py --gpus 2 --steps 500 --warmup 20 --parity-steps 0 --e2e-steps 3 > gpurun_out/r2_bench_n2_s500.json 2> gpurun_out/r2_bench_n2_s500.err; echo "n2 s500 exit $?"
timeout 900 $T bench.py --gpus 2 --hr --strong --nx 3536 --steps 300 --warmup 20 --e2e-steps 2 > gpurun_out/r2_bench_configD_n2.json 2> gpurun_out/r2_bench_configD_n2.err; echo "configD n2 exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench_n2*.json')+glob.glob('gpurun_out/r2_bench_configD*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=d['checks']['parity']
        print(f, '%.4f ms  %.2f G  frac %.3f'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'parity', (p or {}).get('max_rel'), (p or {}).get('normwise'), (p or {}).get('ok'), d['detail']['per_rank_ms_per_step'], d['detail']['per_rank_wait_us_per_step'], 'e2e %.2f G'%(d['e2e']['value']/1e9))
    except Exception as e: print(f,'ERR',e)
PY
tail -3 gpurun_out/*.err | tail -30
