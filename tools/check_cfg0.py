"""cfg0 (GPT-2 small fp32, bs 1-2): the cooperative decode kernel against the launch chain: same tokens, time per generate."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
pkg = importlib.import_module("distributed-lms-raft-llm_b200")
from oracle import reference_path as rp

hf = rp.build_gpt2("small", seed=0)
for B, T, new in ((1, 32, 64), (2, 32, 64), (1, 200, 300), (2, 500, 400)):
    ids = rp.gpt2_prompt_ids(B, T, seed=1)
    outs, times = [], []
    for flag in ("0", "1"):
        os.environ["TUTOR_F32_PERSISTENT"] = flag
        eng = pkg.B200GPT2LMHeadModel.from_hf(hf, dtype="fp32")
        gen = lambda: eng.generate(ids, max_new_tokens=new, repetition_penalty=1.2)
        for _ in range(2):
            out = gen()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(5):
            gen()
        torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / 5 * 1e3
        pre = lambda: eng.generate(ids, max_new_tokens=1, repetition_penalty=1.2)
        pre(); torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(5):
            pre()
        torch.cuda.synchronize(); ms_pre = (time.perf_counter() - t0) / 5 * 1e3
        outs.append(out); times.append((ms, ms_pre))
        del eng
    same = torch.equal(outs[0], outs[1])
    print(f"B={B} T={T} new={new}: chain {times[0][0]:.2f} ms ({B*new/times[0][0]*1e3:.0f} tok/s, fwd {(times[0][0]-times[0][1])/(new-1)*1e3:.1f} us)"
          f"  cooperative {times[1][0]:.2f} ms ({B*new/times[1][0]*1e3:.0f} tok/s, fwd {(times[1][0]-times[1][1])/(new-1)*1e3:.1f} us)  tokens equal: {same}", flush=True)
    if not same:
        d = (outs[0] != outs[1]).nonzero()
        print("  first difference at", d[0].tolist())
