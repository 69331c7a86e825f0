#!/bin/bash
# N = 8, 4, 2 bench lines on one 8-GPU box (usage: gpurun --gpus 8 -- bash tools/gpu_multi.sh TAG)
TAG=${1:-multi}
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29530+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_n$n.json 2> gpurun_out/${TAG}_bench_n$n.err
done
python - $TAG <<'PY'
import json,sys
tag=sys.argv[1]
for n in (2,4,8):
    d=json.load(open("gpurun_out/%s_bench_n%d.json"%(tag,n))); m=d["multi_gpu"]
    print(n, "weak %.4g %.3f ms | replicas %.3f ms | bcast %.3f gather %.3f | strong %.4g %.3f ms (no coll %.3f) | e2e %.4g" % (d["value"], d["ms_per_step"], m["replicas_only"]["ms_per_step"], m["collectives_alone"]["broadcast_ms"], m["collectives_alone"]["all_gather_ms"], m["strong"]["value"], m["strong"]["ms_per_step"], m["strong"]["ms_per_step_without_collectives"], d["e2e"]["value"]))
PY
