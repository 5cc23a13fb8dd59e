#!/bin/bash
# Live contribution of each kernel type to the configs[2] decode forward: leave it out of the step and time the rest.
mkdir -p gpurun_out
: > gpurun_out/ablate.txt
one() { desc="$1"; shift; out=$(env "$@" timeout 300 python bench.py --steps 2 --warmup 3 --no-extras 2>/dev/null | grep '^{' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms/forward' % j['decode']['ms_per_forward'])"); echo "$desc: $out" | tee -a gpurun_out/ablate.txt; }
one "full step" A=1
one "without attention" TUTOR_SKIP=a
one "without c_attn" TUTOR_SKIP=q
one "without c_proj" TUTOR_SKIP=p
one "without c_fc" TUTOR_SKIP=f
one "without mlp c_proj" TUTOR_SKIP=m
one "without the 4 GEMMs" TUTOR_SKIP=qpfm
one "without LM head + finalize" TUTOR_SKIP=h
one "attention only (no GEMMs, no LM head)" TUTOR_SKIP=qpfmh
one "nothing but embed/stats/ln_f" TUTOR_SKIP=qapfmh
