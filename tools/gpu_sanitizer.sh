#!/bin/bash
# compute-sanitizer memcheck + racecheck over __graft_entry__.smoke() (one small hot-path call checked
# against the oracle). Run on the GPU box:  gpurun --timeout 400 -- 'bash tools/gpu_sanitizer.sh r02'
# Not yet run in round 1 (the pod had no free slot on the two attempts at the end of the round).
tag=${1:-rXX}
mkdir -p gpurun_out
for tool in memcheck racecheck; do
  timeout 180 compute-sanitizer --tool $tool --error-exitcode 9 \
    python -c 'import __graft_entry__ as g; g.smoke(); print("SMOKE OK")' \
    > gpurun_out/${tag}_${tool}.log 2>&1
  echo "$tool rc=$?"; tail -4 gpurun_out/${tag}_${tool}.log
done
