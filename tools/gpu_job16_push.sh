#!/bin/bash
# usage: gpu_job16_push.sh N   (reciprocal-force push: tests at N=2, A/B bench at any N)
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
python -m pytest tests/test_mgpu.py "tests/test_plugin.py::test_several_gpus_in_one_context_through_the_plugin" -m gpu -q 2>&1 | tail -25 > gpurun_out/r2_tests16_2gpu.log
tail -3 gpurun_out/r2_tests16_2gpu.log
fi
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N $ARGS > gpurun_out/r2_push${N}_$tag.json 2> gpurun_out/r2_push${N}_$tag.err; }
ARGS="--steps 100 --warmup 40 --no-plugin" run nopush B200_RECIP_PUSH=0 B200_PEER_TRACE=1
ARGS="--steps 100 --warmup 40 --no-plugin" run push B200_RECIP_PUSH=1 B200_PEER_TRACE=1
ARGS="--steps 100 --warmup 40" run full X=1
python - <<PY
import json
for k in ("nopush", "push", "full"):
    try:
        d=json.loads(open(f"gpurun_out/r2_push${N}_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["ms_per_step"],3), "bit", d["bit_identical_to_one_rank"], "plugin", d.get("plugin") and (d["plugin"]["ms_per_step"], d["plugin"]["e2e"] and d["plugin"]["e2e"]["ms_per_step"]), "refgpu", d.get("reference_gpu") and d["reference_gpu"]["ms_per_step"])
    except Exception as e: print(k, "ERR", e)
PY
grep "step time line" gpurun_out/r2_push${N}_nopush.err | tail -$N
grep "step time line" gpurun_out/r2_push${N}_push.err | tail -$N
