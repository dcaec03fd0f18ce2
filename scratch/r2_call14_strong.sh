#!/bin/bash
# config E strong scaling: the SAME 100 M-cell mesh (7072 x 7072 x 2) on N GPUs; usage: r2_call14_strong.sh N
cd /root/repo; mkdir -p gpurun_out
N=$1
if [ "$N" = 1 ]; then
  timeout 1500 python bench.py --strong --nx 7072 --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 1 --parity-steps 2 > gpurun_out/r2_configE_strong_n1.json 2> gpurun_out/r2_configE_strong_n1.err
else
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --strong --nx 7072 --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 1 --parity-steps 2 > gpurun_out/r2_configE_strong_n$N.json 2> gpurun_out/r2_configE_strong_n$N.err
fi
echo "exit $?"; tail -3 gpurun_out/r2_configE_strong_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_configE_strong_n$N.json').read().strip().splitlines()[-1]); p=d['checks']['parity'] or {}
print('N=%d %.4f ms %.1f G frac %.3f'%(d['n_gpus'], d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'parity', p.get('max_rel'), p.get('ok'), d['detail']['per_rank_ms_per_step'], d['detail']['setup_s'])
PY
free -g | head -2
