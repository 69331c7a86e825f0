#!/bin/bash
# Round 2, call B: parity tests + the rewritten bench line (parity sample, free workload, clearance, executed work, CPU arm)
TAG=${1:-r02b}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -5 gpurun_out/${TAG}_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -3 gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err
cat gpurun_out/${TAG}_bench.json | cut -c1-6000
cat gpurun_out/${TAG}_ref.json | cut -c1-1500
nproc; lscpu | grep "Model name"
