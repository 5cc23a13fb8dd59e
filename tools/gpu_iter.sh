#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run k_lm  python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -q -x -k "lmhead or static"
run gpt2_b16 python -m pytest tests/test_gpu_gpt2.py -q -x -s -k "bf16 or cfg2 or cfg3 or padded or launch_modes or ragged or fused"
one() { desc="$1"; shift; out=$(env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-extras 2>/dev/null | grep '^{' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms/forward  %.0f tok/s' % (j['decode']['ms_per_forward'], j['value']))"); echo "$desc: $out"; }
one "default" A=1
one "default" A=1
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_corpus.csv python tools/profile_corpus.py > gpurun_out/ncu_corpus.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_corpus.csv
