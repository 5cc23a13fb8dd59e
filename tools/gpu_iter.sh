#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run k_tc  python -m pytest tests/test_gpu_kernels.py -q -x -k "linear_bf16"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTUTOR_TRACE -I distributed-lms-raft-llm_b200/csrc tools/microbench_gemm.cu -o /tmp/mbg -L distributed-lms-raft-llm_b200 -ltutor_b200 -lcuda 2>/dev/null && LD_LIBRARY_PATH=distributed-lms-raft-llm_b200 timeout 120 /tmp/mbg 2>&1 | head -5 | cut -c1-100
run bert python -m pytest tests/test_gpu_bert.py -q -x -s
