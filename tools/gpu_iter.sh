#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run gpt2_f32 python -m pytest tests/test_gpu_gpt2.py -q -s -k "cfg0 or as_written or tiny or errors or concurrent or launch_modes"
TUTOR_NO_PDL=1 run gpt2_f32_nopdl python -m pytest tests/test_gpu_gpt2.py -q -s -k "cfg0_tokens"
TUTOR_NO_LN_GEMV=1 run gpt2_f32_noln python -m pytest tests/test_gpu_gpt2.py -q -s -k "cfg0_tokens"
