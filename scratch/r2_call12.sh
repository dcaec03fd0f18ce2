#!/bin/bash
# round 2, GPU call 12 (1 GPU): parity suite on the trimmed arithmetic, bench lines (kept), ncu --set full of FO and HR K4 summarised ON THE BOX
cd /root/repo; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_regime.py tests/test_gpu_muscl.py tests/test_reference_decks.py tests/test_gpu_rdycore_adapter.py -m gpu -q --timeout 600 > gpurun_out/r2_pytest_gpu_c12.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r2_pytest_gpu_c12.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_reference_arm.json 2>/dev/null; echo "ref exit $?"
timeout 600 python bench.py --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_s300.json 2>/dev/null; echo "s300 exit $?"
timeout 600 python bench.py --eta0 60 --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_fully_wet.json 2>/dev/null; echo "wet exit $?"
timeout 600 python bench.py --hr --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_configD_hr.json 2>/dev/null; echo "hr exit $?"
timeout 600 python bench.py --hr --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_configD_hr_s20.json 2>/dev/null; echo "hr s20 exit $?"
RDYB_K4_PRIM=0 timeout 600 python bench.py --hr --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_ab4_hr_pm0_s20.json 2>/dev/null; echo "hr pm0 exit $?"
timeout 600 python bench.py --second-order --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2_bench_n1_second_order.json 2>/dev/null; echo "so exit $?"
timeout 600 python bench.py --rk4 --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 --parity-steps 0 > gpurun_out/r2_bench_n1_rk4.json 2>/dev/null; echo "rk4 exit $?"
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
for f in sorted(glob.glob('gpurun_out/r2_bench_n1*.json')+glob.glob('gpurun_out/r2_ab4*.json')+glob.glob('gpurun_out/r2_bench_ref*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=(d.get('checks') or {}).get('parity') or {}
        print(f.split('/')[-1], '%.4f ms  %.3f G  frac %s  parity max_rel %s ok %s'%(d['ms_per_step'], d['value']/1e9, (d.get('roofline') or {}).get('frac'), p.get('max_rel'), p.get('ok')), (d.get('clocks') or {}).get('reasons'))
    except Exception as e: print(f,'ERR',e)
PY
du -sh gpurun_out
