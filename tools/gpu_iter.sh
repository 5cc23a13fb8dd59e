#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-600} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run relev python -m pytest tests/test_gpu_relevance.py -q -x -s
TUTOR_SCAN_SIMT=1 run relev_simt python -m pytest tests/test_gpu_relevance.py -q -x -s
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none --csv --log-file gpurun_out/launches_corpus.csv python tools/profile_corpus.py > gpurun_out/ncu_corpus.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_corpus.csv')) if len(r)>5]
h=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
H=rows[h]
for r in rows[h+1:]:
    d=dict(zip(H,r)); print(d['Kernel Name'][:40], d['Grid Size'], d['Metric Name'], d['Metric Value'], d['Metric Unit'])
PY
