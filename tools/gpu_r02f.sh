#!/bin/bash
# Round 2, call F: executed-work counters of the coarse variant, A/B of TG_COARSE_MEMBERS, ncu of the new phase 1
TAG=${1:-r02f}
mkdir -p gpurun_out
S=$PWD/tygrit_b200/lib/libtygrit_b200_stats.so
for wl in uniform free; do TYGRIT_B200_LIB=$S python tools/executed_work.py --workload $wl | tee -a gpurun_out/${TAG}_executed.jsonl; done
Q="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e"
for v in "" members; do
  L=$PWD/tygrit_b200/lib/libtygrit_b200${v:+_$v}.so
  TYGRIT_B200_LIB=$L $Q > gpurun_out/${TAG}_bench_${v:-base}.json 2> gpurun_out/${TAG}_bench_${v:-base}.err
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_${v:-base}.json")); r=d["roofline"]; f=d.get("free_workload",{})
print("${v:-base}: value %.4g ms/step %.3f p1 %.4f p2 %.4f | free: %.4g cfg/s %.3f ms" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], f.get("value",0), f.get("ms_per_step",0)))
PY
done
EXTRA=l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_sectors.sum,lts__t_sectors.sum,lts__t_requests.sum,lts__t_sectors_srcunit_tex.sum,l1tex__t_requests_pipe_tex.sum,l1tex__t_sectors_pipe_tex.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__inst_executed_op_shared_ld.sum
CMD="python bench.py --configs 33554432 --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:'k_check_configs' -s 3 -c 1 -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
CMDF="python bench.py --configs 33554432 --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras --workload free"
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:'k_check_configs' -s 5 -c 1 -o gpurun_out/${TAG}_prof_free $CMDF > gpurun_out/${TAG}_ncu_full_free.log 2>&1
ls -la gpurun_out | grep ${TAG}
