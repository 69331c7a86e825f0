#!/bin/bash
TAG=${1:-r02r}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log; tail -4 gpurun_out/${TAG}_pytest_gpu.log | cut -c1-300
for v in "" edgefree e256 e1024; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  echo "== ${v:-sync512}"
  TYGRIT_B200_LIB=$L python tools/bench_configs.py --only C2,C5 --no-cpu 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'uniform' in d: print('C2 uniform %.4f ms rrt_like %.4f ms' % (d['uniform']['ms'], d['rrt_like']['ms']))
    else: print('C5 %.3f ms/step build %.0f us' % (d['ms_per_step'], d['build_us']))"
done
