#!/bin/bash
mkdir -p gpurun_out
python tools/debug_plugin.py > gpurun_out/r2_debug_plugin.log 2>&1
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2_tests7.log
tail -4 gpurun_out/r2_tests7.log
python tools/tile_probe.py --tag reverted > gpurun_out/r2_probe7.log 2>&1
tail -1 gpurun_out/r2_probe7.log | cut -c1-400
