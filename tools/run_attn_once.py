"""A few launches of the prefill attention at BERT shape (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
lib, check = lms_b200._lib.load(), lms_b200._lib.check
B, H, T, causal = (int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (32, 12, 512, 0)))
g = torch.Generator(device="cuda").manual_seed(0)
q = torch.randn(B, T, H, 64, device="cuda", generator=g).bfloat16()
k = torch.randn(B, H, T, 64, device="cuda", generator=g).bfloat16()
v = torch.randn(B, H, T, 64, device="cuda", generator=g).bfloat16()
o = torch.empty_like(q)
for _ in range(3):
    check(lib.tutor_attn_prefill(1, q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), B, H, T, T, causal, None))
torch.cuda.synchronize()
import time
for reps in (1, 1, 5, 20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        check(lib.tutor_attn_prefill(1, q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), B, H, T, T, causal, None))
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"{reps} launches: {e0.elapsed_time(e1) * 1e3 / reps:.1f} us per launch on the device, host {1e6 * (t1 - t0) / reps:.1f} us per call")
