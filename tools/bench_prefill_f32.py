"""fp32 prefill of short prompts (the 9..64-row weight-streaming GEMV): GPT-2 small, B x T tokens, CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
hf = rp.build_gpt2("small", seed=0)
eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="fp32")
stream = torch.cuda.ExternalStream(eng.stream)
for B, T in ((1, 32), (1, 64), (2, 24), (1, 9)):
    ids = rp.gpt2_prompt_ids(B, T, seed=1)
    for _ in range(3):
        eng.prefill(ids, max_new_tokens=8)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record(stream)
    for _ in range(n):
        eng.prefill(ids, max_new_tokens=8)
    e1.record(stream)
    torch.cuda.synchronize()
    print(f"fp32 prefill {B} x {T}: {e0.elapsed_time(e1) / n:.3f} ms (host upload included)", flush=True)
