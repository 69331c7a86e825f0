#!/bin/bash
# quick bench (no e2e/cpu) + narrow-phase work counters from the -DTG_STATS build.  usage: tools/gpu_stats.sh <tag> [bench args]
TAG=${1:-s}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json")); r=d["roofline"]; w=d["world"]
print("value %.4g ms/step %.2f p1 %.4f ms p2share %.3f items/cfg %.4f" % (d["value"], d["ms_per_step"], r["avg_launch_ms"], r["phase2_share_of_step"], w["narrow_items_per_config"]))
PY
TYGRIT_B200_LIB=$PWD/tygrit_b200/lib/libtygrit_b200_stats.so python tools/narrow_stats.py "$@" | tee gpurun_out/${TAG}_narrow_stats.json
