#!/bin/bash
TAG=${1:-r02t}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log; tail -4 gpurun_out/${TAG}_pytest_gpu.log | cut -c1-300
for v in "" linear; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  TYGRIT_B200_LIB=$L python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/${TAG}_bench_${v:-tex}.json 2> gpurun_out/${TAG}_bench_${v:-tex}.err
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_${v:-tex}.json")); r=d["roofline"]; f=d.get("free_workload",{})
print("${v:-tex}: value %.4g ms/step %.3f p1 %.4f p2 %.4f build %.3f | free: %.4g cfg/s | clearance %s" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["world"]["build_ms"], f.get("value",0), d.get("with_clearance")))
PY
done
for v in "" linear; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  echo "== ${v:-tex}"
  TYGRIT_B200_LIB=$L python tools/bench_configs.py --only C1,C2,C5 --no-cpu 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'uniform' in d: print('C2 uniform %.4f ms rrt_like %.4f ms' % (d['uniform']['ms'], d['rrt_like']['ms']))
    elif 'device_call_us' in d: print('C1 %.1f us device, %.1f us host' % (d['device_call_us'], d['host_call_us']))
    else: print('C5 %.3f ms/step build %.0f us' % (d['ms_per_step'], d['build_us']))"
done
