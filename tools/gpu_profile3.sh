#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_bf16_persistent_tma -c 8 -f -o gpurun_out/prof_gemm_tma python tools/profile_bert.py > gpurun_out/ncu5.log 2>&1; echo "gemm tma exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"cosine_scan_tf32x3|topk_from_groups" -s 4 -c 2 -f -o gpurun_out/prof_scan_tc2 python tools/profile_corpus.py > gpurun_out/ncu6.log 2>&1; echo "scan exit $?"
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemv_f32_reg -c 6 -f -o gpurun_out/prof_gemv_reg python tools/profile_step.py 64 1 fp32 small > gpurun_out/ncu7.log 2>&1; echo "gemv exit $?"
