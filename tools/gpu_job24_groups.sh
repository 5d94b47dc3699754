#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_bonded.py tests/test_gpu_graph.py "tests/test_reference_suites.py" tests/test_plugin.py -m gpu -q 2>&1 | tail -8
cat gpurun_out/reference_suite_minimizer.json | head -c 700
