#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r2_probe13.log; : > $out
for u in 2 4 6 8; do B200_UNIT_TILES_RT=$u python tools/tile_probe.py --tag stmv_ut$u >> $out 2>&1; done
B200_UNIT_TILES_RT=4 B200_SPARSE_MAX=8 python tools/tile_probe.py --tag stmv_ut4_sp8 >> $out 2>&1
grep "^{" $out | cut -c1-300
python -m pytest tests/test_gpu_methods.py -m gpu -q 2>&1 | tail -3
