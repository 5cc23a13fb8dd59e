#!/bin/bash
# round-2 ncu evidence: launch lists (cold-cache, serialised: shares) and --set full captures of the top kernels
mkdir -p gpurun_out
P="--clock-control none"
timeout 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum $P --csv --log-file gpurun_out/r2c_launches_decode_step.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1; echo "decode launch list exit $?"
timeout 600 ncu --profile-from-start off --set full $P --import-source on -k regex:attn_decode -c 2 -o gpurun_out/r2c_attn_decode python tools/profile_step.py > gpurun_out/ncu2.log 2>&1; echo "attn decode full exit $?"
timeout 300 python tools/profile_step.py 96 1 fp32 > gpurun_out/plain0.log 2>&1 &&
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum $P --csv --log-file gpurun_out/r2c_launches_cfg0_step.csv python tools/profile_step.py 96 1 fp32 > gpurun_out/ncu3.log 2>&1; echo "cfg0 launch list exit $?"
timeout 600 ncu --profile-from-start off --set full $P --import-source on -k regex:decode_f32_persistent -c 1 -o gpurun_out/r2c_cooperative_fp32 python tools/profile_step.py 96 1 fp32 > gpurun_out/ncu4.log 2>&1; echo "cooperative full exit $?"
timeout 300 python tools/profile_bert.py > gpurun_out/plainb.log 2>&1 &&
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum $P --csv --log-file gpurun_out/r2c_launches_bert_32x512.csv python tools/profile_bert.py > gpurun_out/ncu5.log 2>&1; echo "bert launch list exit $?"
timeout 600 ncu --set full $P --import-source on -k regex:attn_prefill_tc -s 2 -c 1 -o gpurun_out/r2c_attn_prefill_tc python tools/run_attn_once.py > gpurun_out/ncu6.log 2>&1; echo "attn prefill full exit $?"
ls -la gpurun_out | tail -20
