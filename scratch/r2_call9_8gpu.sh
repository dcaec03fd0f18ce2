#!/bin/bash
# round 2, the 8-GPU session: 4- and 8-rank parity tests, weak scaling N=1,2,4,8 (20 and 500 steps, per-rank times and in-kernel waits),
# fused push vs push kernel, config E (100 M cells strong, 8 GPUs), config D at 4 GPUs, PCIe ceiling
cd /root/repo; mkdir -p gpurun_out
run() { # name nproc args...
  local name=$1 np=$2; shift 2
  if [ "$np" = 1 ]; then CUDA_VISIBLE_DEVICES=0 timeout 400 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $np "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; fi
  echo "$name exit $?"
}
timeout 420 python -m pytest tests/test_multi_gpu.py -m gpu -q --timeout 300 -k "eight or four" > gpurun_out/r2_pytest_gpu_multi8.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r2_pytest_gpu_multi8.log
run r2_scale_n8_s20 8 --steps 20 --warmup 5 --no-cpu-baseline
run r2_scale_n8_s500 8 --steps 500 --warmup 20 --no-cpu-baseline --parity-steps 0 --e2e-steps 2
RDYB_P2P_NO_FUSED_PUSH=1 run r2_scale_n8_s20_pushkernel 8 --steps 20 --warmup 5 --no-cpu-baseline --parity-steps 0 --e2e-steps 1
run r2_scale_n4_s20 4 --steps 20 --warmup 5 --no-cpu-baseline
run r2_scale_n2_s20 2 --steps 20 --warmup 5 --no-cpu-baseline
run r2_scale_n1_s20 1 --steps 20 --warmup 5 --no-cpu-baseline
run r2_scale_n1_s500 1 --steps 500 --warmup 20 --no-cpu-baseline --parity-steps 0 --e2e-steps 2
run r2_configE_strong_n8_s20 8 --strong --nx 7072 --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2
run r2_configE_strong_n8_s300 8 --strong --nx 7072 --steps 300 --warmup 20 --no-cpu-baseline --parity-steps 0 --e2e-steps 1
run r2_configD_n4 4 --hr --strong --nx 3536 --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29655 tools/pcie_ceiling.py > gpurun_out/r2_pcie_ceiling_n8.json 2>/dev/null; echo "pcie8 exit $?"
CUDA_VISIBLE_DEVICES=0 timeout 120 python tools/pcie_ceiling.py > gpurun_out/r2_pcie_ceiling_n1.json 2>/dev/null; echo "pcie1 exit $?"
nvidia-smi topo -m > gpurun_out/r2_topo.txt 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_scale_*.json')+glob.glob('gpurun_out/r2_config[DE]_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=d['checks']['parity'] or {}
        print(f.split('/')[-1], 'N=%d %.4f ms %.1f G frac %.3f'%(d['n_gpus'], d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'parity', p.get('max_rel'), p.get('ok'), 'ranks', d['detail']['per_rank_ms_per_step'], 'wait', d['detail']['per_rank_mean_wait_us'], d['detail']['per_rank_longest_wait_us'], 'e2e %.2f G %.1f ms'%(d['e2e']['value']/1e9, d['e2e']['ms_per_step']), d['clocks']['reasons'])
    except Exception as e: print(f,'ERR',e)
for f in glob.glob('gpurun_out/r2_pcie_*.json'): print(f, open(f).read()[:600])
PY
