"""configs[0] only: GPT-2 small fp32, B=1, 32-token prompt -> 64 new tokens, reference kwargs; tokens/s over 20 runs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
hf = rp.build_gpt2("small", seed=0)
ids = rp.gpt2_prompt_ids(1, 32, seed=1)
for env in ({}, {"TUTOR_NO_LN_GEMV": "1"}, {"TUTOR_NO_PDL": "1"}, {"TUTOR_NO_LN_GEMV": "1", "TUTOR_NO_PDL": "1"}):
    for k in ("TUTOR_NO_LN_GEMV", "TUTOR_NO_PDL"):
        os.environ.pop(k, None)
    os.environ.update(env)
    eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="fp32")
    for _ in range(5):
        eng.generate(ids, max_new_tokens=64, repetition_penalty=1.2)
    t0 = time.perf_counter()
    n = 20
    for _ in range(n):
        eng.generate(ids, max_new_tokens=64, repetition_penalty=1.2)
    dt = (time.perf_counter() - t0) / n
    print(f"{env or 'default'}: {64 / dt:.0f} tokens/s, {dt * 1e3:.2f} ms per generate", flush=True)
    del eng
