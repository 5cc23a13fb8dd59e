#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout 1200 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-12} gpurun_out/$name.log; }
run new python -m pytest tests/test_gpu_gpt2.py tests/test_gpu_bert.py -m gpu -q -k "ragged"
