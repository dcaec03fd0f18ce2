#!/bin/bash
# round 2, GPU call 5 (2 GPUs): driver (device-side loop), adapter, bench-regime + HR tests; HR bench; config D at 2 GPUs again (HR race fixed)
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_rdy_driver.py tests/test_gpu_rdycore_adapter.py tests/test_gpu_bench_regime.py tests/test_reference_decks.py tests/test_mms.py -m gpu -q --timeout 600 > gpurun_out/r2_pytest_gpu_b.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/r2_pytest_gpu_b.log
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q --timeout 600 -k "hydrostatic or many_tiles" > gpurun_out/r2_pytest_gpu_multi2_hr.log 2>&1; echo "pytest multi hr exit $?"; tail -5 gpurun_out/r2_pytest_gpu_multi2_hr.log
timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_hr.json 2> gpurun_out/r2_bench_n1_hr.err; echo "hr exit $?"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
timeout 900 $T bench.py --gpus 2 --hr --strong --nx 3536 --steps 300 --warmup 20 --e2e-steps 2 > gpurun_out/r2_bench_configD_n2.json 2> gpurun_out/r2_bench_configD_n2.err; echo "configD n2 exit $?"
timeout 900 $T bench.py --gpus 2 --hr --strong --nx 3536 --steps 20 --warmup 5 --e2e-steps 2 --halo nccl > gpurun_out/r2_bench_configD_n2_nccl.json 2> gpurun_out/r2_bench_configD_n2_nccl.err; echo "configD n2 nccl exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench_n1_hr.json')+glob.glob('gpurun_out/r2_bench_configD*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=d['checks']['parity']
        print(f, '%.4f ms  %.2f G  frac %.3f'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'parity', (p or {}).get('max_rel'), (p or {}).get('normwise'), (p or {}).get('ok'), d['detail']['per_rank_ms_per_step'], d['detail']['per_rank_mean_wait_us'], 'e2e %.2f G'%(d['e2e']['value']/1e9))
    except Exception as e: print(f,'ERR',e)
PY
