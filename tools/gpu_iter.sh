#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run k_f32 python -m pytest tests/test_gpu_kernels.py -q -x -k "linear_f32"
run gpt2_f32 python -m pytest tests/test_gpu_gpt2.py -q -k "cfg0 or as_written or tiny or errors or concurrent or launch_modes"
python tools/probe_cfg0.py 2>&1 | grep -E "us/step|generate" | tail -2
TUTOR_GEMV_PLAIN=1 python tools/probe_cfg0.py 2>&1 | grep -E "us/step|generate" | tail -2
