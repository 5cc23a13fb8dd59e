"""Persistent decode kernel (TUTOR_MEGA=2 whole step, 1 per-layer launches) against the 5-launch chain (0) and the fp32
oracle: decode logits on the tiny model over batch sizes that select every row-tile width, then step time on configs[2]."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200
from oracle import reference_path as rp

MODES = sys.argv[1].split(",") if len(sys.argv) > 1 else ["0", "1"]
what = sys.argv[2] if len(sys.argv) > 2 else "tiny,small"


def engine(hf, mode):
    os.environ["TUTOR_MEGA"] = mode
    eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="bf16")
    os.environ.pop("TUTOR_MEGA")
    return eng


if "tiny" in what:
    tiny = rp.build_gpt2("tiny", seed=3)
    for B, T, steps in ((1, 9, 6), (5, 9, 12), (16, 70, 5), (33, 40, 7), (64, 130, 9), (100, 20, 4)):
        ids = rp.gpt2_prompt_ids(B, T, seed=11 + B, vocab=1000)
        res = {}
        for mode in MODES:
            eng = engine(tiny, mode)
            eng.prefill(ids, max_new_tokens=steps + 4, eos_token_id=999, pad_token_id=999)
            eng.step(steps)
            res[mode] = (eng.last_logits().float().cpu(), eng.read_ids())
            del eng
        # fp32 oracle logits for the same (teacher-forced on mode-0 ids so every mode is compared at the same context)
        base = res[MODES[0]]
        with torch.no_grad():
            ref = tiny(base[1][:, :-1]).logits[:, -1].float()
        for mode in MODES:
            lg, out = res[mode]
            same_ctx = torch.equal(out[:, :-1], base[1][:, :-1])
            err = ((lg - ref).abs().max() / ref.abs().max()).item()
            dif = ((lg - base[0]).abs().max() / base[0].abs().max()).item()
            print(f"tiny B={B:3d} T={T:3d} mode {mode}: rel err vs fp32 oracle {err:.3e} (same context: {same_ctx}), vs mode {MODES[0]} {dif:.3e}, "
                  f"ids equal to mode {MODES[0]}: {torch.equal(out, base[1])}", flush=True)

if "small" in what:
    hf = rp.build_gpt2("small", seed=0)
    ids = rp.gpt2_prompt_ids(64, 512, seed=1)
    outs = {}
    for mode in MODES:
        eng = engine(hf, mode)
        stream = torch.cuda.ExternalStream(eng.stream)
        eng.prefill(ids, max_new_tokens=256)
        eng.step(10)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        eng.step(200)
        e1.record(stream)
        torch.cuda.synchronize()
        lg = eng.last_logits().float().cpu()
        outs[mode] = (lg, eng.read_ids())
        print(f"small B=64 mode {mode}: {e0.elapsed_time(e1) * 5:.1f} us per decode forward", flush=True)
        t0 = time.time()
        a = eng.generate(ids, max_new_tokens=256)
        b = eng.generate(ids, max_new_tokens=256)
        print(f"   generate x2 {time.time() - t0:.2f} s, deterministic: {torch.equal(a, b)}", flush=True)
        del eng
        torch.cuda.empty_cache()
    base = outs[MODES[0]]
    for mode in MODES[1:]:
        lg, out = outs[mode]
        agree = (out == base[1]).float().mean().item()
        print(f"small mode {mode} vs {MODES[0]}: token agreement {agree:.4f}")

if "trace" in what:
    import ctypes as C
    import numpy as np
    hf = rp.build_gpt2("small", seed=0)
    ids = rp.gpt2_prompt_ids(64, 512, seed=1)
    for mode in MODES:
        if mode == "0":
            continue
        eng = engine(hf, mode)
        eng.prefill(ids, max_new_tokens=256)
        eng.step(4)
        for cta in (0, 1):
            lms_b200._lib.check(min(0, eng.lib.tutor_gpt2_decode_trace(eng._h, cta, None, 0)))
            eng.step(3)
            buf = np.zeros(128 * 8 + 64 * 8 + 8, dtype=np.int64)
            eng.lib.tutor_gpt2_decode_trace(eng._h, -1, buf.ctypes.data_as(C.c_void_p), buf.size)
            t0, t1 = buf[-1], buf[-2]
            print(f"--- mode {mode} cta {cta}: kernel body {t1 - t0} cycles (last launch of the step in mode 1)")
            st = buf[:1024].reshape(128, 8)
            for i in range(128):
                if st[i, 2] == 0:
                    continue
                row = " ".join(f"{(v - t0) if v else -1:8d}" for v in st[i, [0, 2, 6, 3, 4, 1, 7, 5]])
                print(f"step {i:3d}: {row}")
            un = buf[1024:1024 + 512].reshape(64, 8)
            for i in range(64):
                if un[i, 0] == 0:
                    continue
                row = " ".join(f"{(v - t0) if v else -1:8d}" for v in un[i, :5])
                print(f"unit {i:3d} (q={un[i, 5]}): {row}")
        del eng
