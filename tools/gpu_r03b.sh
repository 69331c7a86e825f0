#!/bin/bash
# A/B: base pose of the next tile prefetched during the current one (k_check_configs)
Q="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e"
for v in "" pf "" pf; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  TYGRIT_B200_LIB=$L $Q > gpurun_out/r03b_bench_${v:-base}.json 2> gpurun_out/r03b_bench_${v:-base}.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r03b_bench_${v:-base}.json")); r=d["roofline"]; f=d.get("free_workload",{})
print("${v:-base}: value %.4g ms/step %.3f p1 %.4f p2 %.4f | free: %.4g cfg/s %.3f ms | mism %d" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], f.get("value",0), f.get("ms_per_step",0), d["parity_sample"]["mismatches"]))
PY
done
