#!/bin/bash
# Round 2, call A (tree as of round-1 HEAD): sanitizers + the ncu coverage VERDICT r01 asked for.
#   1. compute-sanitizer memcheck / racecheck / initcheck over smoke(), memcheck + racecheck over one GPU test per
#      kernel family (queue overflow, fixed-base edges, streaming self-filter, ingest, occlusion)
#   2. ncu --set full (+ the L1->XBAR request counter and L2 sector / request counts) of k_check_configs / k_resolve
#      on the bench launch shape, of k_check_edges (C2) and of the K1 build kernels (C5), k_self_filter_depth (C3 stride 4)
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_r02a.sh r02a'
TAG=${1:-r02a}
mkdir -p gpurun_out
EXTRA=l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_sectors.sum,lts__t_sectors.sum,lts__t_requests.sum,lts__t_sectors_srcunit_tex.sum,l1tex__t_requests_pipe_tex.sum,l1tex__t_sectors_pipe_tex.sum

SMOKE='import __graft_entry__ as g; g.smoke(); print("SMOKE OK")'
for tool in memcheck racecheck initcheck; do
  timeout 300 compute-sanitizer --tool $tool --error-exitcode 9 python -c "$SMOKE" > gpurun_out/${TAG}_sanitizer_${tool}_smoke.log 2>&1
  echo "$tool smoke rc=$?"; tail -3 gpurun_out/${TAG}_sanitizer_${tool}_smoke.log
done
TESTS="tests/test_gpu_parity.py::test_queue_overflow_resolves_inline_with_same_verdicts tests/test_gpu_parity.py::test_self_filter_fast_path_equals_generic_kernel tests/test_gpu_ingest.py tests/test_gpu_callers.py"
for tool in memcheck racecheck; do
  timeout 420 compute-sanitizer --tool $tool --error-exitcode 9 python -m pytest $TESTS -x -q -m gpu > gpurun_out/${TAG}_sanitizer_${tool}_tests.log 2>&1
  echo "$tool tests rc=$?"; tail -3 gpurun_out/${TAG}_sanitizer_${tool}_tests.log
done

CMD="python bench.py --configs 33554432 --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:'k_check_configs|k_resolve' -s 6 -c 2 -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
C2="python tools/bench_configs.py --only C2"
$C2 > gpurun_out/${TAG}_c2_plain.log 2>&1 &&
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:'k_check_edges' -s 3 -c 1 -o gpurun_out/${TAG}_prof_edges $C2 > gpurun_out/${TAG}_ncu_full_edges.log 2>&1
C5="python tools/bench_configs.py --only C5 --steps 4 --vpm 72"
$C5 > gpurun_out/${TAG}_c5_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_count|k_scatter|k_edt_xy|k_edt_z|k_scan' -s 14 -c 7 -o gpurun_out/${TAG}_prof_k1 $C5 > gpurun_out/${TAG}_ncu_full_k1.log 2>&1
C3="python tools/bench_configs.py --only C3 --states 256"
$C3 > gpurun_out/${TAG}_c3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_self_filter_depth' -s 3 -c 1 -o gpurun_out/${TAG}_prof_k4d $C3 > gpurun_out/${TAG}_ncu_full_k4d.log 2>&1
ls -la gpurun_out | tail -30
