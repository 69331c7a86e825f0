#!/bin/bash
# 2 GPUs: NCCL test with the real checker, the multi-GPU bench line, NCCL-touch localisation
TAG=${1:-r02o}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_planner.py -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu_multi.log; tail -4 gpurun_out/${TAG}_pytest_gpu_multi.log | cut -c1-300
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_n2.json 2> gpurun_out/${TAG}_bench_n2.err; echo "bench2 exit $?"; tail -3 gpurun_out/${TAG}_bench_n2.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_n2.json")); m=d["multi_gpu"]; r=d["roofline"]
print("N=2 weak %.4g (%.3f ms) p1 %.4f p2 %.4f | replicas %.3f ms | strong %.4g (%.3f ms, w/o coll %.3f) | bcast %.3f ms allgather %.3f ms | e2e %.4g" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], m["replicas_only"]["ms_per_step"], m["strong"]["value"], m["strong"]["ms_per_step"], m["strong"]["ms_per_step_without_collectives"], m["collectives_alone"]["broadcast_ms"], m["collectives_alone"]["all_gather_ms"], d["e2e"]["value"]))
PY
bash tools/gpu_r02k.sh ${TAG}
