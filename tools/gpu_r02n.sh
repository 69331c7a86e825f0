#!/bin/bash
TAG=${1:-r02n}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log; tail -6 gpurun_out/${TAG}_pytest_gpu.log | cut -c1-300
python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('bench value %.4g with_clearance %s' % (d['value'], d.get('with_clearance')))"
