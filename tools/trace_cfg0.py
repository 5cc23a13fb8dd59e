"""Per-barrier timeline (clock64 of thread 0 of one CTA) of the cooperative fp32 decode kernel on configs[0]."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import lms_b200
from oracle import reference_path as rp

hf = rp.build_gpt2("small", seed=0)
eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="fp32")
ids = rp.gpt2_prompt_ids(1, 32, seed=1)
eng.prefill(ids, max_new_tokens=64, repetition_penalty=1.2)
eng.step(8)
names = ["embed"] + [f"L{l}.{n}" for l in range(12) for n in ("qkv", "attn", "proj", "fc", "proj2")]
for cta in (0, 147):
    eng.lib.tutor_gpt2_decode_trace(eng._h, cta, None, 0)
    eng.step(4)
    buf = np.zeros(1024, dtype=np.int64)
    eng.lib.tutor_gpt2_decode_trace(eng._h, -1, buf.ctypes.data_as(C.c_void_p), 1024)
    t0 = buf[0]
    print(f"CTA {cta}: cycles since the kernel's start; per barrier: thread 0 done | CTA done | release issued | all arrived")
    prev = t0
    tot = {"compute": 0, "skew": 0, "release": 0, "wait": 0}
    for i, n in enumerate(names):
        a, b, c, d = buf[4 * i + 1:4 * i + 5]
        if i < 11 or i > 55:
            print(f"  {n:10s} compute {a - prev:6d}  cta skew {b - a:6d}  release {c - b:6d}  wait {d - c:6d}   (t = {d - t0})")
        tot["compute"] += a - prev; tot["skew"] += b - a; tot["release"] += c - b; tot["wait"] += d - c
        prev = d
    print("  totals:", {k: int(v) for k, v in tot.items()}, "kernel cycles", int(prev - t0))
