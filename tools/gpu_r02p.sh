#!/bin/bash
TAG=${1:-r02p}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log; tail -5 gpurun_out/${TAG}_pytest_gpu.log | cut -c1-300
python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json")); r=d["roofline"]; f=d.get("free_workload",{})
print("early-compact: value %.4g ms/step %.3f p1 %.4f p2 %.4f | free: %.4g cfg/s %.3f ms" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], f.get("value",0), f.get("ms_per_step",0)))
print(r.get("executed"))
PY
