#!/bin/bash
TAG=${1:-r02i}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_planner.py -m gpu -x -q 2>&1 | tail -25
timeout 300 python tools/bench_planner.py > gpurun_out/${TAG}_bench_planner.json 2> gpurun_out/${TAG}_bench_planner.err; echo "planner bench exit $?"; tail -3 gpurun_out/${TAG}_bench_planner.err; cat gpurun_out/${TAG}_bench_planner.json
