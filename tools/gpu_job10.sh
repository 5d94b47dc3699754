#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2_tests10.log
tail -3 gpurun_out/r2_tests10.log
out=gpurun_out/r2_probe10.log; : > $out
python tools/tile_probe.py --tag stmv_pme_flat >> $out 2>&1
python tools/tile_probe.py --workload dhfr --steps 100 --tag dhfr >> $out 2>&1
python tools/tile_probe.py --workload apoa1 --steps 100 --tag apoa1 >> $out 2>&1
grep "^{" $out | cut -c1-330
