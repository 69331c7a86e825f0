#!/bin/bash
# coarse far-field grid size (bytes of shared memory): 128 KB / 176 KB (default) / 186 KB
Q="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extras"
line() { python - "$1" "$2" <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); r=d["roofline"]
print("%-22s value %.4g ms/step %.3f p1 %.4f p2 %.4f mism %d" % (sys.argv[2], d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["parity_sample"]["mismatches"]))
PY
}
for v in "" cg186 cg128; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  for wl in uniform free; do
    TYGRIT_B200_LIB=$L $Q --workload $wl > gpurun_out/r03f_${v:-base}_$wl.json 2>gpurun_out/r03f_${v:-base}_$wl.err && line gpurun_out/r03f_${v:-base}_$wl.json "${v:-base} $wl" || tail -2 gpurun_out/r03f_${v:-base}_$wl.err
  done
done
