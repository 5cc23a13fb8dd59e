#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-4} gpurun_out/$name.log; }
run all python -m pytest tests -m gpu -q -s -x
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1
grep "^{" gpurun_out/bench.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'decode ms', round(d['decode']['ms_per_forward'],4), 'prefill ms', round(d['decode']['prefill_ms'],2), 'frac', round(d['decode']['frac_of_hbm_peak'],3), 'attn roofline', d['roofline']['frac'], d['roofline']['us_per_launch']); print(d.get('extras'))"
timeout 600 python tools/profile_bert.py > gpurun_out/plain_bert.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bert.csv python tools/profile_bert.py > gpurun_out/ncu_bert.log 2>&1
echo "bert launch list exit $?"
