#!/bin/bash
# edge kernel: warp votes between the stages of a round (A1 | A2 | B), register caps
for v in "" e7 e6; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  echo "== ${v:-votes}"
  TYGRIT_B200_LIB=$L python tools/bench_configs.py --only C2,C5 --no-cpu 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'uniform' in d: print('C2 uniform %.4f ms rrt_like %.4f ms' % (d['uniform']['ms'], d['rrt_like']['ms']))
    else: print('C5 %.3f ms/step build %.0f us' % (d['ms_per_step'], d['build_us']))"
done
