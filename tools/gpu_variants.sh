#!/bin/bash
# A/B of library variants built with `python -m tygrit_b200.build --variant NAME -D...`.  usage: tools/gpu_variants.sh "<bench args>" NAME...
ARGS="$1"; shift
mkdir -p gpurun_out
for v in "" "$@"; do
  LIB=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  TYGRIT_B200_LIB=$LIB python bench.py --configs 16777216 --steps 3 --warmup 3 --no-e2e --no-cpu $ARGS > gpurun_out/var.json 2> gpurun_out/var.err || tail -3 gpurun_out/var.err
  python - "${v:-default}" <<'PY'
import json,sys
d=json.load(open("gpurun_out/var.json")); r=d["roofline"]; w=d["world"]
print("%-12s value %.4g ms/step %.3f p1 %.4f ms p2 %.4f ms items/cfg %.4f" % (sys.argv[1], d["value"], d["ms_per_step"], r["avg_launch_ms"],
   r["phase2_share_of_step"]*d["ms_per_step"]/(d["config"]["configs_per_gpu_per_step"]/r["configs_per_launch"]), w["narrow_items_per_config"]))
PY
done
