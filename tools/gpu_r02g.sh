#!/bin/bash
TAG=${1:-r02g}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json")); r=d["roofline"]; f=d.get("free_workload",{})
print("N=1 value %.4g ms/step %.3f p1 %.4f p2 %.4f | free: %.4g cfg/s %.3f ms | clr %.4g" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], f.get("value",0), f.get("ms_per_step",0), d.get("with_clearance",{}).get("value",0)))
PY
for v in 64 80 90; do python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extras --vpm $v > gpurun_out/${TAG}_bench_vpm$v.json 2>/dev/null; python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_vpm$v.json")); r=d["roofline"]
print("vpm $v: value %.4g ms/step %.3f p1 %.4f p2 %.4f build %.3f items %.4f" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["world"]["build_ms"], d["world"]["narrow_items_per_config"]))
PY
done
