#!/bin/bash
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTUTOR_TRACE -I distributed-lms-raft-llm_b200/csrc tools/microbench.cu -o gpurun_out/microbench -L distributed-lms-raft-llm_b200 -ltutor_b200 -Xlinker -rpath=$PWD/distributed-lms-raft-llm_b200 2>&1 | grep -E "error" | head
timeout 120 gpurun_out/microbench 2>&1 | grep -E "stream|flat|attn" | tee gpurun_out/microbench_stream.txt
rm -f gpurun_out/microbench
