#!/bin/bash
# A/B of library variants on the secondary configs (edges).  usage: tools/gpu_variants_c2.sh NAME...
for v in "" "$@" "" "$@"; do
  LIB=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  TYGRIT_B200_LIB=$LIB python tools/bench_configs.py --only C2,C5 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l)
    if 'uniform' in d: print('%-10s C2 uniform %.4f ms rrt %.4f ms' % ('${v:-default}', d['uniform']['ms'], d['rrt_like']['ms']))
    elif 'ms_per_step' in d: print('%-10s C5 %.3f ms/step' % ('${v:-default}', d['ms_per_step']))
"
done
