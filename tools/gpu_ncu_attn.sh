#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_mma -s 20 -c 1 -f -o gpurun_out/prof_attn_mma python tools/profile_attn.py > gpurun_out/ncu_attn_mma.log 2>&1; echo "mma exit $?"
TUTOR_ATTN_SIMT=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_tma -s 20 -c 1 -f -o gpurun_out/prof_attn_simt python tools/profile_attn.py > gpurun_out/ncu_attn_simt.log 2>&1; echo "simt exit $?"
ls -la gpurun_out/*.ncu-rep
