#!/bin/bash
TAG=${1:-r02j}
mkdir -p gpurun_out
python tools/debug_fcit.py 2>&1 | grep -v DEBUG | tail -6
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25
python tools/bench_configs.py --only C2,C5 --no-cpu > gpurun_out/${TAG}_bench_configs.jsonl 2> gpurun_out/${TAG}_bench_configs.err; echo "configs exit $?"; tail -3 gpurun_out/${TAG}_bench_configs.err
python - <<PY
import json
for l in open("gpurun_out/${TAG}_bench_configs.jsonl"):
    d=json.loads(l)
    if "uniform" in d: print("C2 uniform %.4f ms rrt_like %.4f ms" % (d["uniform"]["ms"], d["rrt_like"]["ms"]))
    else: print("C5 %.3f ms/step build %.0f us python loop %.3f" % (d["ms_per_step"], d["build_us"], d["python_loop_ms_per_step"]))
PY
