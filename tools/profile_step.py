"""One decode forward of configs[2] (GPT-2 small bf16, B=64, cache length ~640) inside a cudaProfiler range, for
   ncu --profile-from-start off.  Usage: python tools/profile_step.py [S_target] [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200
from oracle import reference_path as rp

S = int(sys.argv[1]) if len(sys.argv) > 1 else 640
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
dtype = sys.argv[3] if len(sys.argv) > 3 else "bf16"
arch = sys.argv[4] if len(sys.argv) > 4 else "small"
T = 512 if S > 520 else max(1, S - 8)
hf = rp.build_gpt2(arch, seed=0)
eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype=dtype)
ids = rp.gpt2_prompt_ids(B, T, seed=1)
eng.prefill(ids, max_new_tokens=S - T + 8)
eng.step(S - T - 2)          # eager first step, then graph replays
torch.cuda.synchronize()
torch.cuda.profiler.start()
eng.step(1)                  # the profiled forward (one graph launch)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled one decode forward at cache length", eng.read_ids().shape[1] - 1, "launches", lms_b200._lib.load().tutor_launch_count())
