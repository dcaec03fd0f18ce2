#!/bin/bash
# usage: scratch/sweep.sh "<kernel:tile> ..." [extra bench args]
for kt in $1; do
  k=${kt%%:*}; t=${kt##*:}
  timeout 180 python bench.py --kernel $k --tile $t --steps 200 --warmup 10 --no-cpu-baseline --e2e-steps 2 ${@:2} 2>&1 | python -c "
import sys,json
txt=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(txt[-1]); print('$k', d['config']['tile_cells'], 'ms %.4f'%d['ms_per_step'], 'Gcells/s %.2f'%(d['value']/1e9), 'frac %.3f'%d['roofline']['frac'], 'mhz', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'finite', d['checks']['state_finite'])
except Exception as e:
    print('FAILED', '$kt', txt[-5:])
"
done
