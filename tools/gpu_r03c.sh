#!/bin/bash
# k_resolve re-tune at 90 voxels/m: block size / blocks per SM (built variants) and the narrow-phase cell shape (runtime)
Q="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extras"
line() { python - "$1" "$2" <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); r=d["roofline"]
print("%-22s value %.4g ms/step %.3f p1 %.4f p2 %.4f mism %d" % (sys.argv[2], d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["parity_sample"]["mismatches"]))
PY
}
for v in "" r128 r64 r96m8 r96m12; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  for wl in uniform free; do
    TYGRIT_B200_LIB=$L $Q --workload $wl > gpurun_out/r03c_${v:-base}_$wl.json 2>/dev/null && line gpurun_out/r03c_${v:-base}_$wl.json "${v:-base} $wl"
  done
done
for shape in "64 10" "64 12" "64 16" "64 20" "48 14" "96 14" "128 14" "96 20"; do
  set -- $shape
  for wl in uniform free; do
    $Q --workload $wl --cpmx $1 --cpm $2 > gpurun_out/r03c_cell_$1_$2_$wl.json 2>/dev/null && line gpurun_out/r03c_cell_$1_$2_$wl.json "cells x$1 yz$2 $wl"
  done
done
