#!/bin/bash
# r02x: block-cooperative clearance with the half-resolution grid: parity tests, then timing of the variants
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "clearance or oracle or attachment or obstacle or self or shape or tabletop or c1" 2>&1 | tail -3
for v in "" b1024 b512g16 b256 lanes; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  TYGRIT_B200_LIB=$L python tools/clearance_time.py 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('%-28s %s  verdict %.3f ms  +clearance %.2f ms  %.3e /s  ratio %.4f  maxerr %.2e bad %d' % (d['lib'], d['workload'], d['verdict_ms'], d['verdict_plus_clearance_ms'], d['configs_per_s_with_clearance'], d['ratio'], d['max_abs_err_m'], d['n_err_gt_3e-6']))"
done
TYGRIT_B200_LIB=$PWD/tygrit_b200/lib/libtygrit_b200.so python tools/clearance_time.py --workload free | tail -1
