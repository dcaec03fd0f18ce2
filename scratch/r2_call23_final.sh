#!/bin/bash
# round 2, call 23: the final tree - whole GPU suite, smoke, default and second-order bench lines, ncu --set full of K4M
cd /root/repo; mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/r2_final_pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -4 gpurun_out/r2_final_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_final_smoke.log 2>&1; echo "smoke exit $?"; tail -5 gpurun_out/r2_final_smoke.log
timeout 500 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_final_bench_s20.json 2> gpurun_out/r2_final_bench_s20.err; echo "bench exit $?"
B="timeout 200 python bench.py --second-order --steps 20 --warmup 5 --no-cpu-baseline"
$B > gpurun_out/r2_final_bench_so_s20.json 2> gpurun_out/r2_final_bench_so_s20.err; echo "bench so exit $?"
k=swe_tiled_muscl_fixed
timeout 400 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o /tmp/prof_$k -f $B --steps 3 --warmup 3 --e2e-steps 1 --parity-steps 0 > gpurun_out/ncu_$k.log 2>&1; echo "ncu $k exit $?"
python scratch/ncu_summary.py /tmp/prof_$k.ncu-rep > gpurun_out/r2_ncu_full_$k.txt 2>&1
ncu -i /tmp/prof_$k.ncu-rep --page source --csv > /tmp/src_$k.csv 2>/dev/null; python scratch/src_hist.py /tmp/src_$k.csv 9999392 1.5 >> gpurun_out/r2_ncu_full_$k.txt 2>&1
ncu -i /tmp/prof_$k.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; d=rows[2]
for k in ('dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','dram__cycles_active.avg.pct_of_peak_sustained_elapsed','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__inst_executed.sum','lts__t_bytes.sum','lts__t_sector_hit_rate.pct','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__occupancy_limit_shared_mem','launch__grid_size'):
    if k in h: print(k, rows[1][h.index(k)], d[h.index(k)])
" >> gpurun_out/r2_ncu_full_$k.txt
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_final_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); p=(d.get('checks') or {}).get('parity') or {}
        print(f.split('/')[-1], '%.4f ms %.2f G frac %.3f'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac']), 'parity', p.get('max_rel'), p.get('ok'), 'e2e %.2f G'%(d['e2e']['value']/1e9), d['clocks'])
    except Exception as e: print(f, 'unreadable', e)
PY
tail -12 gpurun_out/r2_ncu_full_$k.txt
