#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run relev python -m pytest tests/test_gpu_relevance.py tests/test_serving.py -q -x -m gpu
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_corpus.csv python tools/profile_corpus.py > gpurun_out/ncu_corpus.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_corpus.csv
