#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_mgpu.py "tests/test_plugin.py::test_several_gpus_in_one_context_through_the_plugin" -m gpu -q 2>&1 | tail -25 > gpurun_out/r2_tests9_2gpu.log
tail -3 gpurun_out/r2_tests9_2gpu.log
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 $ARGS > gpurun_out/r2_bench2_$tag.json 2> gpurun_out/r2_bench2_$tag.err; }
ARGS="--steps 20 --warmup 5 --no-plugin" run driverlike1 B200_BENCH_VERBOSE=1
ARGS="--steps 20 --warmup 5 --no-plugin" run driverlike2 X=1
ARGS="--steps 60 --warmup 30 --no-plugin" run fusedpub B200_FUSED_PUBLISH=1
ARGS="--steps 60 --warmup 30 --no-plugin" run separate B200_SEPARATE_REDUCE=1
ARGS="--steps 60 --warmup 30 --no-plugin" run ordered X=1
ARGS="--steps 60 --warmup 30" run full X=1
python - <<'PY'
import json
for k in ("driverlike1", "driverlike2", "fusedpub", "separate", "ordered", "full"):
    try:
        d=json.loads(open(f"gpurun_out/r2_bench2_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["ms_per_step"],3), "bit", d["bit_identical_to_one_rank"], "plugin", d.get("plugin") and (d["plugin"]["ms_per_step"], d["plugin"]["e2e"] and d["plugin"]["e2e"]["ms_per_step"]), "refgpu", d.get("reference_gpu") and d["reference_gpu"]["ms_per_step"])
    except Exception as e: print(k, "ERR", e)
PY
grep "single steps" gpurun_out/r2_bench2_driverlike1.err
