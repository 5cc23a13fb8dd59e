"""Decode-forward time, persistent kernel (TUTOR_MEGA=1) vs launch chain (0), over model / batch shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
models = {}
for arch, B, T, new in (("small", 1, 32, 64), ("small", 8, 256, 64), ("small", 16, 512, 64), ("small", 32, 512, 64), ("small", 128, 512, 64), ("large", 16, 896, 128)):
    if arch not in models: models[arch] = rp.build_gpt2(arch, seed=0)
    ids = rp.gpt2_prompt_ids(B, T, seed=1)
    for mode in ("0", "1"):
        os.environ["TUTOR_MEGA"] = mode
        eng = lms_b200.B200GPT2LMHeadModel.from_hf(models[arch], dtype="bf16")
        stream = torch.cuda.ExternalStream(eng.stream)
        eng.prefill(ids, max_new_tokens=new)
        eng.step(5)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); eng.step(50); e1.record(stream); torch.cuda.synchronize()
        print(f"{arch} B={B} T={T} mega={mode}: {e0.elapsed_time(e1) * 20:.1f} us per decode forward", flush=True)
        del eng; torch.cuda.empty_cache()
