#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -60 > gpurun_out/r2_tests8.log
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
ncu --set full --clock-control none --import-source on -k regex:"tileKernel|nbBuild|pmeSpreadChunks|pmeInterpolateChunks" -s 24 -c 12 -f -o /tmp/r02_prof python tools/tile_probe.py --steps 6 --tag ncu > gpurun_out/ncu_run.log 2>&1
ls -la /tmp/r02_prof.ncu-rep
ncu -i /tmp/r02_prof.ncu-rep --page raw --csv > gpurun_out/r02_ncu_raw.csv 2> gpurun_out/ncu_export.err
ncu -i /tmp/r02_prof.ncu-rep --page source --csv --kernel-name regex:tileKernel --launch-skip 0 --launch-count 1 > gpurun_out/r02_ncu_source_tile.csv 2>> gpurun_out/ncu_export.err
ncu -i /tmp/r02_prof.ncu-rep --page source --csv --kernel-name regex:pmeSpreadChunks --launch-count 1 > gpurun_out/r02_ncu_source_spread.csv 2>> gpurun_out/ncu_export.err
ncu -i /tmp/r02_prof.ncu-rep --page source --csv --kernel-name regex:pmeInterpolateChunks --launch-count 1 > gpurun_out/r02_ncu_source_interp.csv 2>> gpurun_out/ncu_export.err
sz=$(stat -c %s /tmp/r02_prof.ncu-rep 2>/dev/null || echo 0)
if [ "$sz" -lt 30000000 ] && [ "$sz" -gt 0 ]; then cp /tmp/r02_prof.ncu-rep gpurun_out/; fi
du -sh gpurun_out
