#!/bin/bash
# First GPU call of round 2 (one B200, ~3 GPU-minutes): what round 1 could not run any more.
#   1. K3M (swe_tiled_muscl_async.cuh, never run on hardware) under a short timeout: parity tests, then an A/B against K2M;
#   2. the launch list and one --set full capture of the second-order step (K2M), for profiles/ and roofline.traffic.
# Usage: gpurun --timeout 400 -- 'bash scratch/gpu_round2_first.sh'
mkdir -p gpurun_out
export RDYB_TEST_UNVERIFIED=1
timeout 120 python -m pytest tests/test_gpu_muscl.py -m gpu -q -k async > gpurun_out/r2_pytest_k3m.log 2>&1; echo "k3m tests exit $? (124 = hang: check the mbarrier byte counts)"; tail -15 gpurun_out/r2_pytest_k3m.log | cut -c1-300
B="python bench.py --second-order --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2"
timeout 90 $B > gpurun_out/r2_bench_second_order_k2m.json 2> gpurun_out/r2_k2m.err; echo "k2m bench exit $?"
for tile in 128 224; do
  RDYB_MUSCL_ASYNC=1 timeout 90 $B --tile $tile > gpurun_out/r2_bench_second_order_k3m_t$tile.json 2> gpurun_out/r2_k3m_t$tile.err; echo "k3m tile $tile bench exit $?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_bench_second_order_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "%.3f ms/step %.2f Gcell/s" % (d["ms_per_step"], d["value"] / 1e9), d["config"]["kernel_info"][:50])
    except Exception as e:
        print(f, "unreadable:", e)
PY
S="python bench.py --second-order --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1"
timeout 200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_muscl_k2m.csv $S > gpurun_out/r2_ncu1.log 2>&1; echo "ncu launches exit $?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:swe_tiled_muscl_kernel -s 4 -c 1 -o gpurun_out/r2_prof_k2m -f $S > gpurun_out/r2_ncu2.log 2>&1; echo "ncu full exit $?"
