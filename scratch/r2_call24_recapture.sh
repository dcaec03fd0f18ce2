#!/bin/bash
# round 2, call 24: re-capture of the first-order and HR bench kernel (device_layout.c changed since the stamped capture), HR bench line
cd /root/repo; mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 --parity-steps 0"
for v in fo hr; do
  X=""; [ $v = hr ] && X="--hr"
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:swe_tiled_fixed -s 4 -c 1 -o /tmp/prof_$v -f $B $X > gpurun_out/ncu_$v.log 2>&1; echo "ncu $v exit $?"
  python scratch/ncu_summary.py /tmp/prof_$v.ncu-rep > gpurun_out/r2_final_ncu_full_swe_tiled_fixed_$v.txt 2>&1
  ncu -i /tmp/prof_$v.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; d=rows[2]
for k in ('dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','dram__cycles_active.avg.pct_of_peak_sustained_elapsed','smsp__inst_executed.sum'):
    if k in h: print(k, rows[1][h.index(k)], d[h.index(k)])
" >> gpurun_out/r2_final_ncu_full_swe_tiled_fixed_$v.txt
  tail -5 gpurun_out/r2_final_ncu_full_swe_tiled_fixed_$v.txt
done
timeout 200 python bench.py --hr --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_final_bench_hr_s20.json 2> gpurun_out/r2_final_bench_hr_s20.err; echo "bench hr exit $?"; tail -c 300 gpurun_out/r2_final_bench_hr_s20.json | head -c 10
python -c "
import json
d=json.loads(open('gpurun_out/r2_final_bench_hr_s20.json').read().strip().splitlines()[-1]); print('hr %.4f ms %.2f G frac %.3f parity %s'%(d['ms_per_step'], d['value']/1e9, d['roofline']['frac'], d['checks']['parity']['max_rel']))"
