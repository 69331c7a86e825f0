#!/bin/bash
# arm chain in two passes (all seven enclosing-sphere probes in flight, then the members): configs + edges
V=${1:-arm2}
TYGRIT_B200_LIB=$PWD/tygrit_b200/lib/libtygrit_b200_$V.so python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
Q="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extras"
line() { python - "$1" "$2" <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); r=d["roofline"]
print("%-22s value %.4g ms/step %.3f p1 %.4f p2 %.4f mism %d" % (sys.argv[2], d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["parity_sample"]["mismatches"]))
PY
}
for v in "" $V; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  for wl in uniform free; do
    TYGRIT_B200_LIB=$L $Q --workload $wl > gpurun_out/r03n_${v:-base}_$wl.json 2>gpurun_out/r03n_${v:-base}_$wl.err && line gpurun_out/r03n_${v:-base}_$wl.json "${v:-base} $wl" || tail -2 gpurun_out/r03n_${v:-base}_$wl.err
  done
  TYGRIT_B200_LIB=$L python tools/bench_configs.py --only C1,C2,C5 --no-cpu 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'uniform' in d: print('C2 uniform %.4f ms rrt_like %.4f ms' % (d['uniform']['ms'], d['rrt_like']['ms']))
    elif 'device_call_us' in d: print('C1 %.1f us' % d['device_call_us'])
    else: print('C5 %.3f ms/step build %.0f us' % (d['ms_per_step'], d['build_us']))"
done
