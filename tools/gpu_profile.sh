#!/bin/bash
mkdir -p gpurun_out
echo "=== noalloc alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -q > gpurun_out/noalloc_alone.log 2>&1; echo "exit $?" | tee -a gpurun_out/noalloc_alone.log; tail -3 gpurun_out/noalloc_alone.log
echo "=== full gpu suite"; timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/gpu_suite.log 2>&1; echo "exit $?" | tee -a gpurun_out/gpu_suite.log; tail -4 gpurun_out/gpu_suite.log
echo "=== profile"
timeout 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"; tail -2 gpurun_out/plain.log
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:attn_decode -c 2 -o gpurun_out/prof_attn python tools/profile_step.py > gpurun_out/ncu2.log 2>&1
echo "attn full exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_bf16 -c 5 -o gpurun_out/prof_gemm python tools/profile_step.py > gpurun_out/ncu3.log 2>&1
echo "gemm full exit $?"
ls -la gpurun_out
