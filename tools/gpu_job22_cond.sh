#!/bin/bash
# 1 GPU: conditional IF node around the rebuild kernels of the captured evaluation
mkdir -p gpurun_out
python -m pytest tests/test_gpu_graph.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -6
for w in dhfr apoa1 cellulose; do
python tools/tile_probe.py --workload $w --steps 400 --tag ${w}_cond 2>&1 | tail -1 | cut -c1-120
B200_NO_GRAPH_COND=1 python tools/tile_probe.py --workload $w --steps 400 --tag ${w}_nocond 2>&1 | tail -1 | cut -c1-120
done
