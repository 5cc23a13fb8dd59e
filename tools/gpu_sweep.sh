#!/bin/bash
# A/B sweep of the decode-GEMM split / batch-tile choices on the real bench (device-resident decode forward time)
mkdir -p gpurun_out
: > gpurun_out/sweep.txt
one() { desc="$1"; shift; out=$(env "$@" timeout 300 python bench.py --steps 2 --warmup 3 --no-extras 2>/dev/null | grep '^{' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms/forward  %.0f tok/s' % (j['decode']['ms_per_forward'], j['value']))"); echo "$desc: $out" | tee -a gpurun_out/sweep.txt; }
BASE="TUTOR_BN_RES=16 TUTOR_KS_FC=2 TUTOR_BN_FC=16 TUTOR_BN_QKV=16"
one "base2 (qkv 16/2)" $BASE TUTOR_KS_QKV=2
one "qkv 16/1" $BASE TUTOR_KS_QKV=1
one "qkv 16/4" $BASE TUTOR_KS_QKV=4
one "qkv 16/2 proj ks3" $BASE TUTOR_KS_QKV=2 TUTOR_KS_PROJ=3
one "qkv 16/2 proj ks2" $BASE TUTOR_KS_QKV=2 TUTOR_KS_PROJ=2
one "qkv 16/2 proj2 ks4" $BASE TUTOR_KS_QKV=2 TUTOR_KS_PROJ2=4
one "qkv 16/2 proj2 ks6" $BASE TUTOR_KS_QKV=2 TUTOR_KS_PROJ2=6
one "qkv 16/2 proj ks3 proj2 ks6" $BASE TUTOR_KS_QKV=2 TUTOR_KS_PROJ=3 TUTOR_KS_PROJ2=6
