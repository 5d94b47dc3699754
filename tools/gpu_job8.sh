#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/r2_tests8.log
tail -3 gpurun_out/r2_tests8.log
out=gpurun_out/r2_probe8.log; : > $out
for w in dhfr apoa1 cellulose; do
  python tools/tile_probe.py --workload $w --steps 100 --tag ${w}_split_default >> $out 2>&1
  B200_NB_SPLIT=1 python tools/tile_probe.py --workload $w --steps 100 --tag ${w}_split1 >> $out 2>&1
done
B200_NB_SPLIT=16 python tools/tile_probe.py --workload dhfr --steps 100 --tag dhfr_split16 >> $out 2>&1
B200_NB_SPLIT=2 python tools/tile_probe.py --workload apoa1 --steps 100 --tag apoa1_split2 >> $out 2>&1
grep "^{" $out | cut -c1-330
python tools/tile_probe.py --steps 6 --tag ncu_plain > gpurun_out/ncu_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tileKernel|nbBuild|pmeSpreadChunks|pmeInterpolateChunks|Pass|langevinMiddleStep" -s 60 -c 34 -f -o gpurun_out/r02_prof python tools/tile_probe.py --steps 6 --tag ncu > gpurun_out/ncu_run.log 2>&1
ls -la gpurun_out/r02_prof.ncu-rep
