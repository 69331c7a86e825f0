#!/bin/bash
# Round 2, call C (2 GPUs): multi-GPU bench with collectives inside the timed step + strong scaling; N=1 quick bench first
TAG=${1:-r02c}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu --no-e2e > gpurun_out/${TAG}_bench1.json 2> gpurun_out/${TAG}_bench1.err; echo "bench1 exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench1.json")); r=d["roofline"]
print("N=1 value %.4g ms/step %.3f p1 %.4f p2 %.4f" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"]))
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/${TAG}_bench2.json 2> gpurun_out/${TAG}_bench2.err; echo "bench2 exit $?"; tail -5 gpurun_out/${TAG}_bench2.err
cat gpurun_out/${TAG}_bench2.json | cut -c1-5000
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/${TAG}_ref2.json 2> gpurun_out/${TAG}_ref2.err; echo "ref2 exit $?"
cut -c1-400 gpurun_out/${TAG}_ref2.json
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
