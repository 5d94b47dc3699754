#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/r2_tests26.log
tail -3 gpurun_out/r2_tests26.log
python __graft_entry__.py smoke > gpurun_out/r2_smoke.log 2>&1; tail -1 gpurun_out/r2_smoke.log
python bench.py --steps 100 --warmup 20 > gpurun_out/r2_last_stmv.json 2> gpurun_out/r2_last_stmv.err
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r2_last_reference_arm.json 2> gpurun_out/r2_last_reference_arm.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2_last_stmv.json").read().strip().splitlines()[-1])
print("stmv", round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "plugin", d.get("plugin") and round(d["plugin"]["value"],1), "launches", d["gpu_launches"], d["clocks"])
r=json.loads(open("gpurun_out/r2_last_reference_arm.json").read().strip().splitlines()[-1]); print("reference arm", r["value"], r["cpu_baseline"]["cores"])
PY
