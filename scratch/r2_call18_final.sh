#!/bin/bash
# round 2, final validation of the tree after the K5/K6 removal and the allocation-free mass entry point:
# the new test alone in a fresh process, the whole GPU suite (the new test last), smoke(), the default bench line and the reference arm
cd /root/repo; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q --timeout 200 > gpurun_out/r2_final_pytest_no_alloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -3 gpurun_out/r2_final_pytest_no_alloc_alone.log
timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/r2_final_pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -5 gpurun_out/r2_final_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_final_smoke.log 2>&1; echo "smoke exit $?"; tail -6 gpurun_out/r2_final_smoke.log
timeout 500 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_final_bench_s20.json 2> gpurun_out/r2_final_bench_s20.err; echo "bench exit $?"; tail -c 1500 gpurun_out/r2_final_bench_s20.json
timeout 300 python bench.py --hr --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_final_bench_hr_s20.json 2> gpurun_out/r2_final_bench_hr_s20.err; echo "bench hr exit $?"
timeout 300 python bench.py --second-order --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_final_bench_so_s20.json 2> gpurun_out/r2_final_bench_so_s20.err; echo "bench so exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_final_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=(d.get('checks') or {}).get('parity') or {}
        print(f.split('/')[-1], '%.4f ms %.2f G frac %.3f'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'parity', p.get('max_rel'), p.get('ok'), 'e2e %.2f G'%(d['e2e']['value']/1e9), d['clocks'])
    except Exception as e: print(f, 'unreadable', e)
PY
