#!/bin/bash
# round 2, GPU call 11 (1 GPU): full GPU suite on the current kernels, bench lines, ncu --set full of the FO and HR K4 + launch list
cd /root/repo; mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -m gpu -q --timeout 600 > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/r2_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/r2_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench exit $?"
timeout 600 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_s300.json 2>/dev/null; echo "s300 exit $?"
timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_hr.json 2>/dev/null; echo "hr exit $?"
timeout 600 python bench.py --hr --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_hr_s20.json 2>/dev/null; echo "hr s20 exit $?"
timeout 600 python bench.py --second-order --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_second_order.json 2>/dev/null; echo "so exit $?"
timeout 600 python bench.py --second-order --tile 224 --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_second_order_t224.json 2>/dev/null; echo "so224 exit $?"
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 --parity-steps 0"
$B > gpurun_out/plain_fo.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:swe_tiled_fixed -s 4 -c 1 -o gpurun_out/r2_prof_fixed_fo -f $B > gpurun_out/ncu_fo.log 2>&1; echo "ncu fo exit $?"
$B --hr > gpurun_out/plain_hr.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:swe_tiled_fixed -s 4 -c 1 -o gpurun_out/r2_prof_fixed_hr -f $B --hr > gpurun_out/ncu_hr.log 2>&1; echo "ncu hr exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_bench_fixed.csv $B > gpurun_out/ncu_l.log 2>&1; echo "ncu launches exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench_n1*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=d['checks']['parity'] or {}
        print(f.split('/')[-1], '%.4f ms  %.2f G  frac %.3f  parity max_rel %s ok %s'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], p.get('max_rel'), p.get('ok')), d['clocks']['reasons'], d['detail']['kernel_info'][:50])
    except Exception as e: print(f,'ERR',e)
PY
