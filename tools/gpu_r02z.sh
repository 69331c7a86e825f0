#!/bin/bash
# Round 2 final pass: GPU tests, bench line + reference arm, secondary configs, clearance, ncu launch list, full captures of
# the top kernels with the L1-miss-path counters (each ncu run only after the same command has exited 0 without ncu).
TAG=${1:-r02z}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
python tools/bench_configs.py > gpurun_out/${TAG}_bench_configs.jsonl 2> gpurun_out/${TAG}_bench_configs.err
python tools/clearance_time.py > gpurun_out/${TAG}_clearance.jsonl 2>&1
python tools/clearance_time.py --workload free >> gpurun_out/${TAG}_clearance.jsonl 2>&1
EXTRA=l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_sectors.sum,lts__t_sectors.sum,lts__t_requests.sum,lts__t_sectors_srcunit_tex.sum,l1tex__t_requests_pipe_tex.sum,l1tex__t_sectors_pipe_tex.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__inst_executed_op_shared_ld.sum
CMD="python bench.py --configs 67108864 --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:'k_check_configs|k_resolve' -s 6 -c 2 -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
CL="python tools/clearance_time.py --configs 1048576 --sample 4096"
$CL > gpurun_out/${TAG}_cl_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_clearance_block' -s 1 -c 1 -o gpurun_out/${TAG}_prof_clear $CL > gpurun_out/${TAG}_ncu_full_clear.log 2>&1
tail -3 gpurun_out/${TAG}_pytest_gpu.log
cat gpurun_out/${TAG}_bench.json
cat gpurun_out/${TAG}_clearance.jsonl
