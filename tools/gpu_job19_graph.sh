#!/bin/bash
# 1 GPU: full tests with the captured evaluation + bonded kernels on the PME stream, small-system probes, bench lines
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/r2_tests19.log
tail -3 gpurun_out/r2_tests19.log
for w in dhfr apoa1; do
for ut in 1 2 3; do B200_UNIT_TILES_RT=$ut python tools/tile_probe.py --workload $w --steps 400 --tag ${w}_ut$ut 2>&1 | tail -1 | cut -c1-420; done
for sp in 3 8; do B200_SPARSE_MAX=$sp python tools/tile_probe.py --workload $w --steps 400 --tag ${w}_sp$sp 2>&1 | tail -1 | cut -c1-420; done
done
python tools/tile_probe.py --workload stmv --steps 60 --tag stmv 2>&1 | tail -1 | cut -c1-420
python bench.py --workload dhfr --steps 2000 --warmup 500 > gpurun_out/r2_g_dhfr.json 2> gpurun_out/r2_g_dhfr.err
python bench.py --workload apoa1 --steps 1000 --warmup 300 > gpurun_out/r2_g_apoa1.json 2> gpurun_out/r2_g_apoa1.err
python - <<'PY'
import json
for k in ("dhfr", "apoa1"):
    try:
        d=json.loads(open(f"gpurun_out/r2_g_{k}.json").read().strip().splitlines()[-1])
        print(k, round(d["value"],1), "ns/day", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "plugin", d.get("plugin") and round(d["plugin"]["value"],1), "refgpu", d.get("reference_gpu") and round(d["reference_gpu"]["value"],1), "launches", d["gpu_launches"])
    except Exception as e: print(k, "ERR", e)
PY
B200_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_dhfr.csv python bench.py --workload dhfr --steps 12 --warmup 6 --no-plugin --no-cpu-baseline > gpurun_out/ncu_run_dhfr.log 2>&1
wc -l gpurun_out/r02_launches_dhfr.csv
