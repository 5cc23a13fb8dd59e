#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-600} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run k_tc  python -m pytest tests/test_gpu_kernels.py -q -x -k "linear_bf16 and tcgen05"
run bert python -m pytest tests/test_gpu_bert.py -q -x -s
run gpt2_b16 python -m pytest tests/test_gpu_gpt2.py -q -x -s -k "bf16 or cfg2 or cfg3 or padded or launch_modes or ragged"
TAILN=1 run bench python bench.py --steps 3 --warmup 3
python - <<'PY'
import json
for l in open("gpurun_out/bench.log"):
    if l.startswith("{"):
        j=json.loads(l); print(j["value"], j["decode"], j["extras"]["cfg1_bert_base_bf16_32pairs"])
PY
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bert.csv python tools/profile_bert.py > gpurun_out/ncu_bert.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_bert.csv
