#!/bin/bash
# Runs the GPU test groups in separate processes (a device-side trap poisons only its own process).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; echo "=== $name"; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -5 gpurun_out/$name.log; }
run k_simt   python -m pytest tests/test_gpu_kernels.py -q -k "not bf16 and not lmhead"
run k_legacy python -m pytest tests/test_gpu_kernels.py -q -k "linear_bf16 and mma_sync"
run k_tc     python -m pytest tests/test_gpu_kernels.py -q -k "linear_bf16 and tcgen05"
run k_lmhead python -m pytest tests/test_gpu_kernels.py -q -k "lmhead"
run k_attn16 python -m pytest tests/test_gpu_kernels.py -q -k "attn"
run gpt2_f32 python -m pytest tests/test_gpu_gpt2.py -q -s -k "cfg0_tokens or cfg0_logits or as_written or tiny or errors or concurrent"
run gpt2_b16 python -m pytest tests/test_gpu_gpt2.py -q -s -k "bf16 or cfg2"
run bert     python -m pytest tests/test_gpu_bert.py -q -s
run relev    python -m pytest tests/test_gpu_relevance.py -q -s
run smoke    python __graft_entry__.py --smoke
run bench    python bench.py --steps 2 --warmup 3
