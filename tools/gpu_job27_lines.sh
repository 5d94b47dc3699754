#!/bin/bash
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_final3_stmv.json 2> gpurun_out/r2_final3_stmv.err
python bench.py --workload cellulose --steps 300 --warmup 60 > gpurun_out/r2_final3_cellulose.json 2> gpurun_out/r2_final3_cellulose.err
python bench.py --workload apoa1ljpme --steps 1000 --warmup 300 > gpurun_out/r2_final3_apoa1ljpme.json 2> gpurun_out/r2_final3_apoa1ljpme.err
python bench.py --workload dhfr --steps 2000 --warmup 500 > gpurun_out/r2_final3_dhfr.json 2> gpurun_out/r2_final3_dhfr.err
python bench.py --workload apoa1 --steps 1000 --warmup 300 > gpurun_out/r2_final3_apoa1.json 2> gpurun_out/r2_final3_apoa1.err
python - <<'PY'
import json
for k in ("stmv", "cellulose", "apoa1", "apoa1ljpme", "dhfr"):
    try:
        d=json.loads(open(f"gpurun_out/r2_final3_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), "ns/day", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "plugin", d.get("plugin") and round(d["plugin"]["value"],1), "refgpu", d.get("reference_gpu") and round(d["reference_gpu"]["value"],1), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"],2), "roof", round(d["roofline"]["frac"],3), round(d["roofline"]["pme_frac"] or 0,3), "launches", d["gpu_launches"])
    except Exception as e: print(k, "ERR", e)
PY
