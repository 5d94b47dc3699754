#!/bin/bash
# 1 GPU: A/B of the conditional node on long runs, then the final validation (tests, smoke, bench lines, launch list)
mkdir -p gpurun_out
for i in 1 2; do for w in dhfr apoa1; do
python tools/tile_probe.py --workload $w --steps 4000 --tag ${w}_plain_$i 2>&1 | tail -1 | cut -c1-110
B200_GRAPH_COND=1 python tools/tile_probe.py --workload $w --steps 4000 --tag ${w}_cond_$i 2>&1 | tail -1 | cut -c1-110
done; done
python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/r2_tests23.log
tail -3 gpurun_out/r2_tests23.log
python __graft_entry__.py smoke > gpurun_out/r2_smoke.log 2>&1; tail -1 gpurun_out/r2_smoke.log
python bench.py > gpurun_out/r2_final2_stmv.json 2> gpurun_out/r2_final2_stmv.err
python bench.py --workload cellulose --steps 300 --warmup 60 > gpurun_out/r2_final2_cellulose.json 2> gpurun_out/r2_final2_cellulose.err
python bench.py --workload apoa1ljpme --steps 1000 --warmup 300 > gpurun_out/r2_final2_apoa1ljpme.json 2> gpurun_out/r2_final2_apoa1ljpme.err
python bench.py --workload dhfr --steps 2000 --warmup 500 > gpurun_out/r2_final2_dhfr.json 2> gpurun_out/r2_final2_dhfr.err
python bench.py --workload apoa1 --steps 1000 --warmup 300 > gpurun_out/r2_final2_apoa1.json 2> gpurun_out/r2_final2_apoa1.err
python - <<'PY'
import json
for k in ("stmv", "cellulose", "apoa1", "apoa1ljpme", "dhfr"):
    try:
        d=json.loads(open(f"gpurun_out/r2_final2_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), "ns/day", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "plugin", d.get("plugin") and round(d["plugin"]["value"],1), "refgpu", d.get("reference_gpu") and round(d["reference_gpu"]["value"],1), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"],2), "roof", round(d["roofline"]["frac"],3), round(d["roofline"]["pme_frac"] or 0,3), "launches", d["gpu_launches"])
    except Exception as e: print(k, "ERR", e)
PY
python bench.py --steps 6 --warmup 3 --no-plugin --no-cpu-baseline > gpurun_out/ncu_plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_launches_stmv.csv python bench.py --steps 6 --warmup 3 --no-plugin --no-cpu-baseline > gpurun_out/ncu_run3.log 2>&1
wc -l gpurun_out/r02_launches_stmv.csv
