#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-4} gpurun_out/$name.log; }
run gpt2 python -m pytest tests/test_gpu_gpt2.py tests/test_gpu_kernels.py -q -s -x
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTUTOR_TRACE -I distributed-lms-raft-llm_b200/csrc tools/microbench.cu -o gpurun_out/microbench -L distributed-lms-raft-llm_b200 -ltutor_b200 -Xlinker -rpath=$PWD/distributed-lms-raft-llm_b200 2>&1 | tail -3
timeout 120 gpurun_out/microbench 2>&1 | grep -v chase | grep -v "sync pairs" | tee gpurun_out/microbench.txt
rm -f gpurun_out/microbench
timeout 600 python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/bench.log 2>&1
grep "^{" gpurun_out/bench.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'decode ms', round(d['decode']['ms_per_forward'],4), 'prefill ms', round(d['decode']['prefill_ms'],2), 'frac', round(d['decode']['frac_of_hbm_peak'],3))"
