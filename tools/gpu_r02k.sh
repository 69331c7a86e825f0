#!/bin/bash
# 2 GPUs: which NCCL use makes phase 1 slower?  (tools/p1_time.py --touch ...)
TAG=${1:-r02k}
mkdir -p gpurun_out
OUT=gpurun_out/${TAG}_localise.jsonl
: > $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
port=29530
for touch in "" blob valid staged blob,valid; do
  port=$((port+1))
  $TR --master-port $port tools/p1_time.py --mode nccl --tag "nccl_touch_${touch:-none}" ${touch:+--touch $touch} 2>/dev/null | grep -o '{.*}' >> $OUT
done
port=$((port+1))
NCCL_NVLS_ENABLE=0 $TR --master-port $port tools/p1_time.py --mode nccl --tag "nccl_nvls0_touch_blob_valid" --touch blob,valid 2>/dev/null | grep -o '{.*}' >> $OUT
port=$((port+1))
NCCL_CUMEM_ENABLE=0 $TR --master-port $port tools/p1_time.py --mode nccl --tag "nccl_cumem0_touch_blob_valid" --touch blob,valid 2>/dev/null | grep -o '{.*}' >> $OUT
python - <<PY
import json,re
for m in re.finditer(r'\{[^{}]*\}', open("$OUT").read()):
    d=json.loads(m.group(0)); print('%-34s rank %d  ms/step %.3f  p1 %.4f  p2 %.4f' % (d['tag'],d['rank'],d['ms_per_step'],d['phase1_ms_per_launch'],d['phase2_ms_per_launch']))
PY
