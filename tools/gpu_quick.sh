#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-4} gpurun_out/$name.log; }
run gpt2 python -m pytest tests/test_gpu_gpt2.py tests/test_gpu_kernels.py tests/test_gpu_bert.py tests/test_zz_no_alloc.py -q -s -x
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1
grep "^{" gpurun_out/bench.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'decode ms', round(d['decode']['ms_per_forward'],4), 'prefill ms', round(d['decode']['prefill_ms'],2), 'frac', round(d['decode']['frac_of_hbm_peak'],3), 'attn roofline', d['roofline']['frac'], d['roofline']['us_per_launch']); print(d.get('extras')); print(d.get('cpu_baseline'))"
