#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_step.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"; python tools/summarize_launches.py gpurun_out/launches_step.csv
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_decode_fused -c 5 -f -o gpurun_out/prof_gemm_fused python tools/profile_step.py > gpurun_out/ncu3.log 2>&1
echo "fused gemm full exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cosine_scan_tf32x3 -s 2 -c 1 -f -o gpurun_out/prof_scan_tc python tools/profile_corpus.py > gpurun_out/ncu4.log 2>&1
echo "scan full exit $?"
