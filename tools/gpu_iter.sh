#!/bin/bash
# one development iteration on the GPU box: new-kernel unit tests, the bf16 engine tests, microbench, a short bench
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-600} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run k_fused  python -m pytest tests/test_gpu_kernels.py -q -x -k "decode_ or row_tile or lmhead"
run gpt2_b16 python -m pytest tests/test_gpu_gpt2.py -q -x -s -k "bf16 or cfg2 or padded or launch_modes"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTUTOR_TRACE -I distributed-lms-raft-llm_b200/csrc tools/microbench_fused.cu -o /tmp/mf -L distributed-lms-raft-llm_b200 -ltutor_b200 -lcuda && LD_LIBRARY_PATH=distributed-lms-raft-llm_b200 timeout 120 /tmp/mf > gpurun_out/microbench_fused.txt 2>&1; echo exit $?; cat gpurun_out/microbench_fused.txt
TAILN=1 run bench    python bench.py --steps 3 --warmup 3
TAILN=1 TUTOR_NO_CLUSTER=1 run bench_nocluster python bench.py --steps 3 --warmup 3
python - <<'PY'
import json
for f in ("bench","bench_nocluster"):
    for l in open(f"gpurun_out/{f}.log"):
        if l.startswith("{"):
            j=json.loads(l); print(f, j["value"], j["decode"])
PY
