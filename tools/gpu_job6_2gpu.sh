#!/bin/bash
mkdir -p gpurun_out
for k in 1 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2950$k bench.py --gpus 2 --steps 20 --warmup 5 --no-plugin > gpurun_out/r2_bench_stmv_2gpu_driverlike_$k.json 2> gpurun_out/r2_bench_stmv_2gpu_driverlike_$k.err
done
B200_BENCH_VERBOSE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29503 bench.py --gpus 2 --steps 20 --warmup 5 --no-plugin > gpurun_out/r2_bench_stmv_2gpu_driverlike_3.json 2> gpurun_out/r2_bench_stmv_2gpu_driverlike_3.err
python - <<'PY'
import json
for k in (1,2,3):
    try:
        d=json.loads(open(f"gpurun_out/r2_bench_stmv_2gpu_driverlike_{k}.json").read().strip().splitlines()[-1]); print(k, d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["bit_identical_to_one_rank"])
    except Exception as e: print(k, "ERR", e)
PY
grep "single steps" gpurun_out/r2_bench_stmv_2gpu_driverlike_3.err
