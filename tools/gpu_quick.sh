#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout 1200 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-3} gpurun_out/$name.log; }
run all python -m pytest tests -m gpu -q -x
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1
grep "^{" gpurun_out/bench.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'decode ms', round(d['decode']['ms_per_forward'],4), 'prefill', round(d['decode']['prefill_ms'],2), 'roof', round(d['roofline']['frac'],3)); print(d['extras']); print(d['cpu_baseline'])"
timeout 600 python tools/profile_bert.py > gpurun_out/plain_bert.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bert.csv python tools/profile_bert.py > gpurun_out/ncu_bert.log 2>&1
echo "bert launch list exit $?"
timeout 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "step launch list exit $?"
