#!/bin/bash
# refresh of the round-2 ncu evidence after the tensor-memory-operand attention and the CTA-pair GEMM
mkdir -p gpurun_out
P="--clock-control none"
timeout 300 python tools/profile_bert.py > gpurun_out/plainb.log 2>&1 || { tail -5 gpurun_out/plainb.log; exit 1; }
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum $P --csv --log-file gpurun_out/r2f_launches_bert_32x512.csv python tools/profile_bert.py > gpurun_out/ncu5.log 2>&1; echo "bert launch list exit $?"
timeout 600 ncu --profile-from-start off --set full $P --import-source on -k regex:gemm_bf16_persistent_tma -c 4 -o gpurun_out/r2f_gemm_pair python tools/profile_bert.py > gpurun_out/ncu8.log 2>&1; echo "gemm pair full exit $?"
timeout 600 ncu --set full $P --import-source on -k regex:attn_prefill_tc -s 2 -c 1 -o gpurun_out/r2f_attn_prefill_tc python tools/run_attn_once.py > gpurun_out/ncu6.log 2>&1; echo "attn prefill full exit $?"
ls -la gpurun_out | grep r2f
