#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run k_tc  python -m pytest tests/test_gpu_kernels.py -q -x -k "linear_bf16"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTUTOR_TRACE -I distributed-lms-raft-llm_b200/csrc tools/microbench_gemm.cu -o /tmp/mbg -L distributed-lms-raft-llm_b200 -ltutor_b200 -lcuda 2>/dev/null && LD_LIBRARY_PATH=distributed-lms-raft-llm_b200 timeout 120 /tmp/mbg 2>&1 | head -11 | cut -c1-100
run relev python -m pytest tests/test_gpu_relevance.py -q -x
run bert python -m pytest tests/test_gpu_bert.py -q -x
run gpt2_b16 python -m pytest tests/test_gpu_gpt2.py -q -x -k "bf16 or cfg2 or cfg3 or padded or launch_modes or ragged"
TAILN=1 run bench python bench.py --steps 3 --warmup 3
python - <<'PY'
import json
for l in open("gpurun_out/bench.log"):
    if l.startswith("{"):
        j=json.loads(l); print(j["value"], j["decode"]["prefill_ms"], j["extras"]["cfg1_bert_base_bf16_32pairs"]["ms_per_batch"], j["extras"]["cfg3_gpt2_large_bf16_bs16"], j["extras"]["cfg4_corpus_1M_x768_fp32_32q_top10"])
PY
