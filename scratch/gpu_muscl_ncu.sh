#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --second-order --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1"
timeout 300 $B > gpurun_out/muscl_plain.json 2> gpurun_out/muscl_plain.err; echo "plain exit $?"
timeout 500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_muscl.csv $B > gpurun_out/ncu_muscl.log 2>&1; echo "ncu exit $?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/launches_muscl.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); mi=hdr.index("Metric Name"); vi=hdr.index("Metric Value")
agg=collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[1:]:
    agg[r[ki][:60]][r[mi]].append(float(r[vi].replace(",","")))
for k,m in agg.items():
    print(k, {a:(len(v), sum(v)/len(v)) for a,v in m.items()})
PY
