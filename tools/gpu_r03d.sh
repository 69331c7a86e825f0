#!/bin/bash
# clearance kernel: block size x configurations per block iteration (barrier granularity), after the full GPU test suite
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
for v in "" c256x32 c128x16 c128x32 c64x8; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  for wl in uniform free; do
  TYGRIT_B200_LIB=$L python tools/clearance_time.py --workload $wl 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('%-28s %-8s verdict %.3f ms  +clearance %.2f ms  %.3e /s  ratio %.4f  maxerr %.2e bad %d' % (d['lib'], d['workload'], d['verdict_ms'], d['verdict_plus_clearance_ms'], d['configs_per_s_with_clearance'], d['ratio'], d['max_abs_err_m'], d['n_err_gt_3e-6']))"
  done
done
python tools/bench_configs.py --only C2,C5 --no-cpu 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'uniform' in d: print('C2 uniform %.4f ms rrt_like %.4f ms' % (d['uniform']['ms'], d['rrt_like']['ms']))
    else: print('C5 %.3f ms/step build %.0f us' % (d['ms_per_step'], d['build_us']))"
