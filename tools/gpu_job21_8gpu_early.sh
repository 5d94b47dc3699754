#!/bin/bash
# 8 GPUs: early FORCES signal on the PME rank + float hand-over (default) against the hand-over switched off
N=8
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N $ARGS > gpurun_out/r2_early${N}_$tag.json 2> gpurun_out/r2_early${N}_$tag.err; }
ARGS="--steps 100 --warmup 40 --no-plugin --no-cpu-baseline" run default B200_PEER_TRACE=1
ARGS="--steps 100 --warmup 40 --no-plugin --no-cpu-baseline" run off B200_RECIP_PUSH=0
python - <<PY
import json
for k in ("default", "off"):
    try:
        d=json.loads(open(f"gpurun_out/r2_early${N}_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["ms_per_step"],3), "bit", d["bit_identical_to_one_rank"])
    except Exception as e: print(k, "ERR", e)
PY
grep "step time line" gpurun_out/r2_early${N}_default.err | tail -$N
