#!/bin/bash
mkdir -p gpurun_out
python tools/debug_exc.py > gpurun_out/r2_debug_exc.log 2>&1
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2_tests2.log
tail -5 gpurun_out/r2_tests2.log
