#!/bin/bash
# 2 GPUs: the case that timed out, then tests and the A/B bench of the reciprocal-force push
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/debug_push.py > gpurun_out/r2_dbg_push_$tag.log 2>&1; echo "== $tag"; grep -v "^\[W\|^W1\|\*\*\*\|OMP_NUM\|^$\|NCCL version" gpurun_out/r2_dbg_push_$tag.log | cut -c1-120 | tail -6; }
run stmv_push DBG_WORKLOAD=stmv DBG_CHUNKS=8,8,8,8,8,8,8,8,8,8,8,8
bash tools/gpu_job16_push.sh 2
