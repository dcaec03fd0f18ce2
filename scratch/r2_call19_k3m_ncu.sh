#!/bin/bash
# round 2, call 19: ncu --set full of the second-order step (K3M + the gradient kernel) at 10 M cells, summarised on the box
cd /root/repo; mkdir -p gpurun_out
B="python bench.py --second-order --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 --parity-steps 0"
$B > gpurun_out/plain_so.log 2>&1; echo "plain exit $?"
for k in swe_tiled_muscl_async muscl_gradient; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o /tmp/prof_$k -f $B > gpurun_out/ncu_$k.log 2>&1; echo "ncu $k exit $?"
  python scratch/ncu_summary.py /tmp/prof_$k.ncu-rep > gpurun_out/r2_ncu_full_$k.txt 2>&1
  ncu -i /tmp/prof_$k.ncu-rep --page source --csv > /tmp/src_$k.csv 2>/dev/null; python scratch/src_hist.py /tmp/src_$k.csv 9999392 1.5 >> gpurun_out/r2_ncu_full_$k.txt 2>&1
  ncu -i /tmp/prof_$k.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; d=rows[2]
for k in ('dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','dram__cycles_active.avg.pct_of_peak_sustained_elapsed','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__inst_executed.sum','lts__t_bytes.sum','lts__t_sector_hit_rate.pct','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__occupancy_limit_shared_mem','launch__grid_size'):
    if k in h: print(k, rows[1][h.index(k)], d[h.index(k)])
" >> gpurun_out/r2_ncu_full_$k.txt
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2_launches_bench_so.csv $B > gpurun_out/ncu_l.log 2>&1; echo "ncu launches exit $?"
tail -45 gpurun_out/r2_ncu_full_swe_tiled_muscl_async.txt; tail -12 gpurun_out/r2_ncu_full_muscl_gradient.txt
