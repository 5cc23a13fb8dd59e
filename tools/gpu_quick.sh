#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -${TAILN:-4} gpurun_out/$name.log; }
run gpt2 python -m pytest tests/test_gpu_gpt2.py tests/test_gpu_kernels.py -q -s -x
TAILN=2 run bench python bench.py --steps 3 --warmup 3 --no-extras
timeout 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_bf16 -c 4 -o gpurun_out/prof_gemm_splitk python tools/profile_step.py > gpurun_out/ncu3.log 2>&1
echo "gemm full exit $?"
TUTOR_NO_PDL=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/bench_nopdl.log 2>&1; tail -1 gpurun_out/bench_nopdl.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('NO_PDL value', d['value'], 'decode ms', d['decode']['ms_per_forward'])"
