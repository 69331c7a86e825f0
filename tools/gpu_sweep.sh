#!/bin/bash
# parameter sweep of the config-check bench (short runs, no e2e / cpu legs).  usage: tools/gpu_sweep.sh "<args1>" "<args2>" ...
mkdir -p gpurun_out
for args in "$@"; do
  python bench.py --configs 16777216 --steps 3 --warmup 3 --no-e2e --no-cpu $args > gpurun_out/sweep.json 2> gpurun_out/sweep.err || tail -3 gpurun_out/sweep.err
  python - "$args" <<'PY'
import json,sys
d=json.load(open("gpurun_out/sweep.json")); r=d["roofline"]; w=d["world"]
print("%-24s value %.4g ms/step %.2f p1 %.3f ms p2 %.3f ms items/cfg %.4f build %.3f ms grid %s" % (sys.argv[1], d["value"], d["ms_per_step"],
   r["avg_launch_ms"], r["phase2_share_of_step"]*d["ms_per_step"]/ (d["config"]["configs_per_gpu_per_step"]/r["configs_per_launch"]), w["narrow_items_per_config"], w["build_ms"], w["grid"]))
PY
done
