#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/profile_attn.py > gpurun_out/plain_attn.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 24 -c 3 -o gpurun_out/prof_attn_tma python tools/profile_attn.py > gpurun_out/ncu_attn.log 2>&1
echo "attn full exit $?"
timeout 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"; tail -1 gpurun_out/bench_ref.log | cut -c1-300
