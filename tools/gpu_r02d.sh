#!/bin/bash
# Round 2, call D (2 GPUs): localise the multi-rank slowdown of phase 1
TAG=${1:-r02d}
mkdir -p gpurun_out
OUT=gpurun_out/${TAG}_localise.jsonl
: > $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
python tools/p1_time.py --tag alone_gpu0 --device 0 2>/dev/null | tail -1 >> $OUT
python tools/p1_time.py --tag alone_gpu1 --device 1 2>/dev/null | tail -1 >> $OUT
(python tools/p1_time.py --tag two_procs_gpu0 --device 0 2>/dev/null | tail -1 >> $OUT) & (python tools/p1_time.py --tag two_procs_gpu1 --device 1 2>/dev/null | tail -1 >> ${OUT}.b); wait; cat ${OUT}.b >> $OUT; rm -f ${OUT}.b
$TR --master-port 29521 tools/p1_time.py --mode none --tag torchrun_no_pg 2>/dev/null | grep '^{' >> $OUT
$TR --master-port 29522 tools/p1_time.py --mode gloo --tag torchrun_gloo 2>/dev/null | grep '^{' >> $OUT
$TR --master-port 29523 tools/p1_time.py --mode nccl --tag torchrun_nccl 2>/dev/null | grep '^{' >> $OUT
$TR --master-port 29524 tools/p1_time.py --mode nccl-destroyed --tag torchrun_nccl_destroyed 2>/dev/null | grep '^{' >> $OUT
NCCL_P2P_DISABLE=1 $TR --master-port 29525 tools/p1_time.py --mode nccl --tag torchrun_nccl_p2p_disabled 2>/dev/null | grep '^{' >> $OUT
OMP_NUM_THREADS=8 $TR --master-port 29526 tools/p1_time.py --mode none --tag torchrun_no_pg_omp8 2>/dev/null | grep '^{' >> $OUT
cat $OUT
