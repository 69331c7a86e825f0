#!/bin/bash
# One gpurun call: GPU parity tests, the bench line (+ reference arm), the secondary configs, the ncu launch list and
# full captures of the top kernels (each ncu run only after the same command has exited 0 without ncu).
# usage: tools/gpu_round.sh <tag>
TAG=${1:-v}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err
python tools/bench_configs.py > gpurun_out/${TAG}_bench_configs.jsonl 2> gpurun_out/${TAG}_bench_configs.err
# launch shape of the bench: 32 Mi configurations per launch pair
CMD="python bench.py --configs 33554432 --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_check_configs|k_resolve' -s 6 -c 2 -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
C3="python tools/bench_configs.py --only C3"
$C3 > gpurun_out/${TAG}_c3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_self_filter_stream' -s 3 -c 1 -o gpurun_out/${TAG}_prof_k4 $C3 > gpurun_out/${TAG}_ncu_full_k4.log 2>&1
tail -3 gpurun_out/${TAG}_pytest_gpu.log
cat gpurun_out/${TAG}_bench.json
