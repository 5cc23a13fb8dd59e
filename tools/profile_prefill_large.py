"""One GPT-2 large bf16 prefill of 16 x 1024 tokens (configs[3]) inside a cudaProfiler range for
ncu --profile-from-start off; without ncu it prints the event-timed prefill."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200
from oracle import reference_path as rp

eng = lms_b200.B200GPT2LMHeadModel.from_hf(rp.build_gpt2("large", seed=0), dtype="bf16")
ids = rp.gpt2_prompt_ids(16, 1023, seed=1)
stream = torch.cuda.ExternalStream(eng.stream)
for _ in range(2):
    eng.prefill(ids, max_new_tokens=1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
e0.record(stream)
eng.prefill(ids, max_new_tokens=1)
e1.record(stream)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"prefill 16 x {ids.shape[1]}: {e0.elapsed_time(e1):.2f} ms")
