#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run gpt2_b16 python -m pytest tests/test_gpu_gpt2.py -q -x -k "bf16 or cfg2 or cfg3 or padded or launch_modes or ragged"
one() { desc="$1"; shift; out=$(env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-extras 2>/dev/null | grep '^{' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms/forward  %.0f tok/s' % (j['decode']['ms_per_forward'], j['value']))"); echo "$desc: $out"; }
one "prefetch" A=1
one "no prefetch" TUTOR_NO_WEIGHT_PREFETCH=1
one "prefetch" A=1
one "no prefetch" TUTOR_NO_WEIGHT_PREFETCH=1
