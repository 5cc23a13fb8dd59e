#!/bin/bash
one() { desc="$1"; shift; out=$(env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-extras 2>/dev/null | grep '^{' | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms/forward  %.0f tok/s' % (j['decode']['ms_per_forward'], j['value']))"); echo "$desc: $out"; }
one "default" A=1
one "without c_fc" TUTOR_SKIP=f
one "default" A=1
