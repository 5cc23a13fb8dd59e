"""Where does a batch-1 fp32 decode step go?  GPU time (CUDA events on the engine stream) against host wall time
for a burst of graph-replayed steps."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
hf = rp.build_gpt2("small", seed=0)
ids = rp.gpt2_prompt_ids(1, 32, seed=1)
eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="fp32")
stream = torch.cuda.ExternalStream(eng.stream)
for rep in range(3):
    eng.prefill(ids, max_new_tokens=120)
    eng.step(4)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(stream)
    eng.step(100)
    t_issue = time.perf_counter() - t0
    e1.record(stream)
    torch.cuda.synchronize()
    t_wall = time.perf_counter() - t0
    print(f"100 steps: GPU {e0.elapsed_time(e1) * 10:.1f} us/step, host issue {t_issue * 1e4:.1f} us/step, wall {t_wall * 1e4:.1f} us/step")
t0 = time.perf_counter()
for _ in range(10):
    eng.generate(ids, max_new_tokens=64, repetition_penalty=1.2)
print(f"generate: {(time.perf_counter() - t0) / 10 * 1e3:.2f} ms per 64 tokens")
t0 = time.perf_counter()
for _ in range(10):
    eng.prefill(ids, max_new_tokens=64)
    torch.cuda.synchronize()
print(f"prefill(32 tokens) alone: {(time.perf_counter() - t0) / 10 * 1e3:.2f} ms")
