"""Persistent decode kernel vs the launch chain on the tiny model over (B, groups) combinations."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
tiny = rp.build_gpt2("tiny", seed=3)
def run(B, T, steps, mode, groups=None):
    os.environ["TUTOR_MEGA"] = mode
    if groups: os.environ["TUTOR_MEGA_GROUPS"] = str(groups)
    else: os.environ.pop("TUTOR_MEGA_GROUPS", None)
    eng = lms_b200.B200GPT2LMHeadModel.from_hf(tiny, dtype="bf16")
    ids = rp.gpt2_prompt_ids(B, T, seed=11 + B, vocab=1000)
    eng.prefill(ids, max_new_tokens=steps + 4, eos_token_id=999, pad_token_id=999)
    eng.step(steps)
    return eng.last_logits().float().cpu()
for B, T in ((64, 130), (64, 20), (33, 40), (32, 20), (48, 20), (100, 20), (128, 20)):
    for steps in (1, 5):
        base = run(B, T, steps, "0")
        for groups in (1, 2, 4):
            lg = run(B, T, steps, "1", groups)
            err = (lg - base).abs().amax(dim=1) / base.abs().max()
            bad = (err > 2e-2).nonzero().flatten().tolist()
            print(f"B={B:3d} T={T:3d} steps={steps} groups={groups}: max rel diff {err.max().item():.3e} bad rows {bad[:12]}{'...' if len(bad) > 12 else ''}", flush=True)
