#!/bin/bash
# round 2, GPU call 15 (1 GPU): the full GPU suite on the final tree, smoke, final bench lines, ncu --set full of the final FO / HR K4 (summarised on the box), K3M tile sweep
cd /root/repo; mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -m gpu -q --timeout 600 > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/r2_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/r2_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench exit $?"
timeout 600 python bench.py --hr --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_configD_hr_s20.json 2>/dev/null; echo "hr s20 exit $?"
for t in 96 160 192; do
  timeout 300 python bench.py --second-order --tile $t --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 1 --parity-steps 0 > gpurun_out/r2_so_tile$t.json 2>/dev/null; echo "so tile $t exit $?"
done
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 --parity-steps 0"
for v in fo hr; do
  X=""; [ $v = hr ] && X="--hr"
  $B $X > gpurun_out/plain_$v.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:swe_tiled_fixed -s 4 -c 1 -o /tmp/prof_$v -f $B $X > gpurun_out/ncu_$v.log 2>&1; echo "ncu $v exit $?"
  python scratch/ncu_summary.py /tmp/prof_$v.ncu-rep > gpurun_out/r2_ncu_full_swe_tiled_fixed_$v.txt 2>&1
  ncu -i /tmp/prof_$v.ncu-rep --page source --csv > /tmp/src_$v.csv 2>/dev/null; python scratch/src_hist.py /tmp/src_$v.csv 9999392 1.0 >> gpurun_out/r2_ncu_full_swe_tiled_fixed_$v.txt 2>&1
  ncu -i /tmp/prof_$v.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; d=rows[2]
for k in ('dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','dram__cycles_active.avg.pct_of_peak_sustained_elapsed','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__inst_executed.sum'):
    if k in h: print(k, rows[1][h.index(k)], d[h.index(k)])
" >> gpurun_out/r2_ncu_full_swe_tiled_fixed_$v.txt
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_bench_fixed.csv $B > gpurun_out/ncu_l.log 2>&1; echo "ncu launches exit $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench_n1*.json')+glob.glob('gpurun_out/r2_so_tile*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=(d.get('checks') or {}).get('parity') or {}
        print(f.split('/')[-1], '%.4f ms  %.3f G  frac %.4f  parity max_rel %s ok %s'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], p.get('max_rel'), p.get('ok')), d['clocks']['reasons'], d['roofline'].get('traffic'))
    except Exception as e: print(f,'ERR',e)
PY
tail -8 gpurun_out/r2_ncu_full_swe_tiled_fixed_fo.txt; tail -8 gpurun_out/r2_ncu_full_swe_tiled_fixed_hr.txt
