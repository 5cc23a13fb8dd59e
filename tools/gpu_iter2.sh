#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/bench_decode_gemms.py > gpurun_out/decode_gemms.txt 2>&1; echo "exit $?"; cat gpurun_out/decode_gemms.txt
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_cluster.csv python tools/profile_step.py > gpurun_out/ncu_cluster.log 2>&1
echo "ncu exit $?"; python tools/summarize_launches.py gpurun_out/launches_cluster.csv
