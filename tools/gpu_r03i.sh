#!/bin/bash
# stage A1 with several configurations per lane (ILP 1 / 2 / 3): parity first, then the bench
#python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
Q="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extras"
line() { python - "$1" "$2" <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); r=d["roofline"]
print("%-22s value %.4g ms/step %.3f p1 %.4f p2 %.4f mism %d" % (sys.argv[2], d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["parity_sample"]["mismatches"]))
PY
}
for v in "" s256 s128 s1024; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  [ -f $L ] || continue
  for wl in uniform free; do
    TYGRIT_B200_LIB=$L $Q --workload $wl > gpurun_out/r03i_${v:-ilp3}_$wl.json 2>gpurun_out/r03i_${v:-ilp3}_$wl.err && line gpurun_out/r03i_${v:-ilp3}_$wl.json "${v:-ilp3} $wl" || tail -2 gpurun_out/r03i_${v:-ilp3}_$wl.err
  done
done
