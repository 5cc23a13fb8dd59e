#!/bin/bash
# end-of-session evidence: full GPU suite, the default bench, launch lists of the three workloads
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/gpu_suite.log 2>&1; echo "suite exit $?"; tail -2 gpurun_out/gpu_suite.log
timeout 900 python bench.py > gpurun_out/bench_full.log 2>&1; echo "bench exit $?"; grep '^{' gpurun_out/bench_full.log > gpurun_out/bench_n1.json
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_step.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_step.csv
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bert.csv python tools/profile_bert.py > gpurun_out/ncu_bert.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_bert.csv
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_corpus.csv python tools/profile_corpus.py > gpurun_out/ncu_corpus.log 2>&1
