#!/bin/bash
# 2 GPUs: float hand-over (pull) variant of the reciprocal forces: tests with it on and off; 1-GPU A/B of graph / bonded stream at STMV size
mkdir -p gpurun_out
python -m pytest tests/test_mgpu.py "tests/test_plugin.py::test_several_gpus_in_one_context_through_the_plugin" -m gpu -q 2>&1 | tail -8 > gpurun_out/r2_tests20_2gpu.log
tail -3 gpurun_out/r2_tests20_2gpu.log
N=2
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N $ARGS > gpurun_out/r2_pull${N}_$tag.json 2> gpurun_out/r2_pull${N}_$tag.err; }
ARGS="--steps 100 --warmup 40 --no-plugin" run off B200_RECIP_PUSH=0
ARGS="--steps 100 --warmup 40 --no-plugin" run on B200_RECIP_PUSH=1
python - <<PY
import json
for k in ("off", "on"):
    try:
        d=json.loads(open(f"gpurun_out/r2_pull${N}_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["ms_per_step"],3), "bit", d["bit_identical_to_one_rank"])
    except Exception as e: print(k, "ERR", e)
PY
export CUDA_VISIBLE_DEVICES=0
for i in 1 2; do
python tools/tile_probe.py --workload stmv --steps 100 --tag stmv_default_$i 2>&1 | tail -1 | cut -c1-300
B200_NO_GRAPH=1 python tools/tile_probe.py --workload stmv --steps 100 --tag stmv_nograph_$i 2>&1 | tail -1 | cut -c1-300
B200_BONDED_MAIN=1 python tools/tile_probe.py --workload stmv --steps 100 --tag stmv_bondedmain_$i 2>&1 | tail -1 | cut -c1-300
B200_NO_GRAPH=1 B200_BONDED_MAIN=1 python tools/tile_probe.py --workload stmv --steps 100 --tag stmv_nograph_bondedmain_$i 2>&1 | tail -1 | cut -c1-300
done
