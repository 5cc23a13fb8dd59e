#!/bin/bash
# full GPU test suite, then the default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/gpu_suite.log 2>&1; echo "suite exit $?" | tee -a gpurun_out/gpu_suite.log; tail -4 gpurun_out/gpu_suite.log
timeout 900 python bench.py > gpurun_out/bench_full.log 2>&1; echo "bench exit $?"; grep '^{' gpurun_out/bench_full.log | cut -c1-1500
