#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-6} gpurun_out/$name.log; }
run serving python -m pytest tests/test_serving.py -q -x -m gpu
TAILN=1 run bench python bench.py --steps 3 --warmup 3
python - <<'PY'
import json
for l in open("gpurun_out/bench.log"):
    if l.startswith("{"):
        j=json.loads(l); print(j["value"], j["e2e"]["value"], j["decode"]); print(json.dumps(j["extras"], indent=1)); print(j["cpu_baseline"])
PY
