#!/bin/bash
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/debug_push.py > gpurun_out/r2_dbg_push_$tag.log 2>&1; echo "== $tag"; grep -v "^\[W\|^W1\|\*\*\*\|OMP_NUM\|^$\|NCCL version" gpurun_out/r2_dbg_push_$tag.log | cut -c1-300 | tail -14; }
run stmv_push DBG_WORKLOAD=stmv DBG_CHUNKS=8,8,8,8,8,8,8,8,8,8,8,8
run stmv_push_unconsumed DBG_WORKLOAD=stmv DBG_UNCONSUMED=1 DBG_CHUNKS=40,8,8,8,8,8,8,8
run stmv_join DBG_WORKLOAD=stmv B200_PUSH_JOIN=1 DBG_UNCONSUMED=1 DBG_CHUNKS=40,8,8,8,8,8,8,8
run dhfr_unconsumed DBG_UNCONSUMED=1 DBG_CHUNKS=40,20,20,20,20
