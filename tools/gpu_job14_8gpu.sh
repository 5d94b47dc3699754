#!/bin/bash
N=8
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N $ARGS > gpurun_out/r2_bench${N}_$tag.json 2> gpurun_out/r2_bench${N}_$tag.err; }
ARGS="--steps 100 --warmup 40 --no-plugin" run ordered B200_ORDERED_REDUCE=1 B200_PEER_TRACE=1
ARGS="--steps 100 --warmup 40" run full B200_PEER_TRACE=1
python - <<PY
import json
for k in ("ordered", "full"):
    try:
        d=json.loads(open(f"gpurun_out/r2_bench${N}_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["ms_per_step"],3), "bit", d["bit_identical_to_one_rank"], "plugin", d.get("plugin") and (d["plugin"]["ms_per_step"], d["plugin"]["e2e"] and d["plugin"]["e2e"]["ms_per_step"]), "refgpu", d.get("reference_gpu") and d["reference_gpu"]["ms_per_step"])
    except Exception as e: print(k, "ERR", e)
PY
grep "step time line" gpurun_out/r2_bench${N}_full.err | tail -8
