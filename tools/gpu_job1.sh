#!/bin/bash
# first GPU job of round 2: parity tests, STMV + DHFR bench lines (with plugin / reference CUDA rows)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_tests1.log
python bench.py --steps 100 --warmup 30 > gpurun_out/r2_bench_stmv_a.json 2> gpurun_out/r2_bench_stmv_a.err
python bench.py --workload dhfr --steps 300 --warmup 100 > gpurun_out/r2_bench_dhfr_a.json 2> gpurun_out/r2_bench_dhfr_a.err
tail -3 gpurun_out/r2_tests1.log
