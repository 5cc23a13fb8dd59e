"""Prefill / bidirectional attention: the tcgen05 kernel against the mma.sync one (TUTOR_ATTN_PREFILL_MMA=1), CUDA events."""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) < 2:
    for flag in ("0", "1"):
        env = dict(os.environ)
        if flag == "1":
            env["TUTOR_ATTN_PREFILL_MMA"] = "1"
        print("mma.sync kernel" if flag == "1" else "tcgen05 kernel", flush=True)
        subprocess.run([sys.executable, __file__, "run"], env=env, check=True)
    sys.exit(0)
import torch
import lms_b200
lib, check = lms_b200._lib.load(), lms_b200._lib.check
dev = "cuda"

def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps

g = torch.Generator(device=dev).manual_seed(0)
for (B, H, T, causal) in ((32, 12, 512, 0), (32, 12, 128, 0), (64, 12, 512, 1), (16, 20, 1024, 1)):
    q = torch.randn(B, T, H, 64, device=dev, generator=g).bfloat16()
    k = torch.randn(B, H, T, 64, device=dev, generator=g).bfloat16()
    v = torch.randn(B, H, T, 64, device=dev, generator=g).bfloat16()
    o = torch.empty_like(q)
    us = timeit(lambda: check(lib.tutor_attn_prefill(1, q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), B, H, T, T, causal, None)))
    fl = 4.0 * T * T * 64 * H * B * (0.5 if causal else 1.0)
    print(f"  attn_prefill bf16 B={B} H={H} T={T} causal={causal}: {us:8.1f} us  {fl / us * 1e-6:7.1f} TFLOP/s", flush=True)
