#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2_tests3.log
tail -3 gpurun_out/r2_tests3.log
out=gpurun_out/r2_probe3.log; : > $out
python tools/tile_probe.py --tag base >> $out 2>&1
for sm in 3 8 12; do B200_SPARSE_MAX=$sm python tools/tile_probe.py --tag sparse$sm >> $out 2>&1; done
for v in w8_b4_u16 w4_b8_u16 w8_b3_u8; do B200_LIB_PATH=build/variants/libb200nb_$v.so python tools/tile_probe.py --tag $v >> $out 2>&1; done
B200_NB_PADDING=0.12 python tools/tile_probe.py --tag pad12 >> $out 2>&1
cat $out | grep -v "^$" | tail -12
