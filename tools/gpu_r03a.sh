#!/bin/bash
# sub-block decoupling of k_check_configs (named barriers): A/B of the sub-block size, parity first
python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
Q="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e"
for v in "" sub1024 sub512 sub128; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  TYGRIT_B200_LIB=$L $Q > gpurun_out/r03a_bench_${v:-sub256}.json 2> gpurun_out/r03a_bench_${v:-sub256}.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r03a_bench_${v:-sub256}.json")); r=d["roofline"]; f=d.get("free_workload",{})
print("${v:-sub256}: value %.4g ms/step %.3f p1 %.4f p2 %.4f | free: %.4g cfg/s %.3f ms | mism %d" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], f.get("value",0), f.get("ms_per_step",0), d["parity_sample"]["mismatches"]))
PY
done
