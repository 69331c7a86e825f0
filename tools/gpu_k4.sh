#!/bin/bash
# K4 A/B: parity tests of the self-filter + the C3 line for each library variant.  usage: tools/gpu_k4.sh NAME...
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "self_filter or filter" 2>&1 | tail -5
for v in "" "$@"; do
  LIB=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  echo "== ${v:-default}"
  TYGRIT_B200_LIB=$LIB python tools/bench_configs.py --only C3 2>&1 | tail -1 | tee -a gpurun_out/k4_variants.jsonl
done
