#!/bin/bash
# 4 GPUs: multi-GPU parity tests (world 2 and 4, hand-over off and on), then the bench line of the final code
mkdir -p gpurun_out
python -m pytest tests/test_mgpu.py -m gpu -q 2>&1 | tail -5 > gpurun_out/r2_tests25_4gpu.log
tail -3 gpurun_out/r2_tests25_4gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 100 --warmup 40 --no-cpu-baseline > gpurun_out/r2_final_4gpu.json 2> gpurun_out/r2_final_4gpu.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2_final_4gpu.json").read().strip().splitlines()[-1])
print(round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["ms_per_step"],3), "bit", d["bit_identical_to_one_rank"], "plugin", d.get("plugin") and d["plugin"]["ms_per_step"], "refgpu", d.get("reference_gpu") and d["reference_gpu"]["ms_per_step"])
PY
