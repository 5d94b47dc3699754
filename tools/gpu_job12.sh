#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/r2_tests12.log
tail -3 gpurun_out/r2_tests12.log
out=gpurun_out/r2_probe12.log; : > $out
for u in 1 2 4 8; do B200_UNIT_TILES_RT=$u python tools/tile_probe.py --workload dhfr --steps 150 --tag dhfr_ut$u >> $out 2>&1; done
for u in 2 4 8; do B200_UNIT_TILES_RT=$u python tools/tile_probe.py --workload apoa1 --steps 100 --tag apoa1_ut$u >> $out 2>&1; done
for u in 4 8; do B200_UNIT_TILES_RT=$u python tools/tile_probe.py --workload cellulose --steps 60 --tag cellulose_ut$u >> $out 2>&1; done
grep "^{" $out | cut -c1-300
