#!/bin/bash
# quick GPU iteration: parity tests + short bench (no ncu).  usage: tools/gpu_quick.sh <tag> [bench args]
TAG=${1:-q}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/${TAG}_pytest_gpu.log
python bench.py --steps 5 --warmup 3 "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_bench.json"))
    r=d["roofline"]; w=d["world"]
    print("value %.4g cfg/s  ms/step %.2f  e2e %.4g  p1 share %.3f p2 share %.3f  p1 launch ms %.4f items/cfg %.4f def %.3f free %.3f build %.3f ms cpu %.4g" % (
      d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value",0), r["share_of_step"] or 0, r["phase2_share_of_step"] or 0, r["avg_launch_ms"],
      w["narrow_items_per_config"], w["definite_collision_fraction"], w["free_fraction"], w["build_ms"], (d.get("cpu_baseline") or {}).get("value",0)))
except Exception as e: print("no bench json", e)
PY
