#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -30 > gpurun_out/r2_tests5.log
tail -3 gpurun_out/r2_tests5.log
out=gpurun_out/r2_probe5.log; : > $out
python tools/tile_probe.py --tag merged >> $out 2>&1
for v in ut12 ut16; do B200_LIB_PATH=build/variants/libb200nb_$v.so python tools/tile_probe.py --tag $v >> $out 2>&1; done
B200_SPARSE_MAX=8 B200_LIB_PATH=build/variants/libb200nb_ut16.so python tools/tile_probe.py --tag ut16_sp8 >> $out 2>&1
python tools/tile_probe.py --workload dhfr --steps 200 --tag dhfr >> $out 2>&1
python tools/cufft_crosscheck.py >> $out 2>&1
grep "^{" $out | cut -c1-420
