"""Bisect run-to-run nondeterminism of bf16 generation (GPT-2 large, B=16, 896 -> +128) over the engine's debug knobs."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200
from oracle import reference_path as rp

arch = sys.argv[1] if len(sys.argv) > 1 else "large"
B, T, NEW = (16, 896, 128) if arch == "large" else (64, 512, 256)
hf = rp.build_gpt2(arch, seed=0)
ids = rp.gpt2_prompt_ids(B, T, seed=1)
for env in ({}, {"TUTOR_NO_PDL": "1"}, {"TUTOR_NO_GRAPH": "1"}, {"TUTOR_NO_FUSED": "1"}, {"TUTOR_NO_PDL": "1", "TUTOR_NO_GRAPH": "1"}):
    for k in ("TUTOR_NO_PDL", "TUTOR_NO_GRAPH", "TUTOR_NO_FUSED"):
        os.environ.pop(k, None)
    os.environ.update(env)
    eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="bf16")
    outs = [eng.generate(ids, max_new_tokens=NEW) for _ in range(4)]
    diffs = [(o != outs[0]).any(dim=1).sum().item() for o in outs[1:]]
    first = [int((o != outs[0]).any(dim=0).float().argmax()) if (o != outs[0]).any() else -1 for o in outs[1:]]
    print(f"{env or 'default'}: rows differing from run 0: {diffs}, first differing position: {first}", flush=True)
    del eng
