#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-600} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run k_attn  python -m pytest tests/test_gpu_kernels.py -q -x -k "attn_decode"
run gpt2_b16 python -m pytest tests/test_gpu_gpt2.py -q -x -s -k "bf16 or cfg2 or cfg3 or padded or launch_modes or ragged"
TUTOR_ATTN_SPLITS=3 run gpt2_b16_s3 python -m pytest tests/test_gpu_gpt2.py -q -x -s -k "bf16 or cfg2 or cfg3 or padded or launch_modes or ragged"
one() { desc="$1"; shift; out=$(env "$@" timeout 300 python bench.py --steps 2 --warmup 3 --no-extras 2>/dev/null | grep '^{' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms/forward  %.0f tok/s' % (j['decode']['ms_per_forward'], j['value']))"); echo "$desc: $out" | tee -a gpurun_out/sweep.txt; }
: > gpurun_out/sweep.txt
one "default" A=1
one "splits=1" TUTOR_ATTN_SPLITS=1
one "splits=2" TUTOR_ATTN_SPLITS=2
one "splits=3" TUTOR_ATTN_SPLITS=3
one "splits=4" TUTOR_ATTN_SPLITS=4
one "simt" TUTOR_ATTN_SIMT=1
