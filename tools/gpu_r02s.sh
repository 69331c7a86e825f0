#!/bin/bash
# N GPUs: the driver's scaling command, as the driver launches it
N=${1:-8}; TAG=${2:-r02s}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench exit $?"; tail -3 gpurun_out/${TAG}_bench_n$N.err | cut -c1-300
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_n$N.json")); m=d["multi_gpu"]; r=d["roofline"]
print("N=$N weak %.4g (%.3f ms) p1 %.4f p2 %.4f | replicas %.3f ms | strong %.4g (%.3f ms, w/o coll %.3f) | bcast %.3f ms (%.0f GB/s) allgather %.3f ms | e2e %.4g" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], m["replicas_only"]["ms_per_step"], m["strong"]["value"], m["strong"]["ms_per_step"], m["strong"]["ms_per_step_without_collectives"], m["collectives_alone"]["broadcast_ms"], m["collectives_alone"]["broadcast_GBps"], m["collectives_alone"]["all_gather_ms"], d["e2e"]["value"]))
print("per-rank ms", [round(x,3) for x in m["ms_per_step_per_rank"]]); print("clocks", [c["sm_mhz"] for c in m["clocks_per_rank"]], [c["reasons"] for c in m["clocks_per_rank"]])
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --impl reference --gpus $N --steps 2 --warmup 1 2>/dev/null | cut -c1-300
