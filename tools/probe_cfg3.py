"""configs[3] decode step time (GPT-2 large bf16, B=16, cache ~960) under the current environment knobs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
hf = rp.build_gpt2("large", seed=0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ids = rp.gpt2_prompt_ids(B, 896, seed=1)
VARIANTS = [{}, {"TUTOR_ATTN_SPLITS": "1"}]
for env in VARIANTS:
    for k in ("TUTOR_ATTN_SPLITS", "TUTOR_ATTN_MMA"):
        os.environ.pop(k, None)
    os.environ.update(env)
    eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="bf16")
    stream = torch.cuda.ExternalStream(eng.stream)
    eng.prefill(ids, max_new_tokens=128)
    eng.step(10)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    eng.step(100)
    e1.record(stream)
    torch.cuda.synchronize()
    print(f"B={B} {env or 'default'}: {e0.elapsed_time(e1) * 10:.1f} us per decode forward", flush=True)
    del eng
    torch.cuda.empty_cache()
