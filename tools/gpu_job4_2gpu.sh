#!/bin/bash
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo.txt 2>&1
python -m pytest tests/test_mgpu.py -m gpu -q 2>&1 | tail -8 > gpurun_out/r2_tests_mgpu2.log
B200_PEER_TRACE=1 B200_BENCH_VERBOSE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus 2 --steps 60 --warmup 30 --no-plugin > gpurun_out/r2_bench_stmv_2gpu_a.json 2> gpurun_out/r2_bench_stmv_2gpu_a.err
tail -3 gpurun_out/r2_tests_mgpu2.log; grep "b200 rank\|rank 0\|rank 1" gpurun_out/r2_bench_stmv_2gpu_a.err | tail -8
