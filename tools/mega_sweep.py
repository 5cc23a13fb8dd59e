"""Decode-forward time of configs[2] (GPT-2 small, B=64, cache 512+) under environment knobs given as K=V,... sets."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
hf = rp.build_gpt2("small", seed=0)
ids = rp.gpt2_prompt_ids(64, 512, seed=1)
for spec in sys.argv[1:]:
    env = dict(kv.split("=") for kv in spec.split(",") if kv)
    os.environ.update(env)
    eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="bf16")
    stream = torch.cuda.ExternalStream(eng.stream)
    eng.prefill(ids, max_new_tokens=256)
    eng.step(10)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    eng.step(200)
    e1.record(stream)
    torch.cuda.synchronize()
    print(f"{spec or 'default'}: {e0.elapsed_time(e1) * 5:.1f} us per decode forward", flush=True)
    for k in env: os.environ.pop(k)
    del eng
    torch.cuda.empty_cache()
