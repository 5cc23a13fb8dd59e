"""Stand-alone timings (CUDA events, 20 reps after warm-up) of the prefill-side kernels at BERT / GPT-2 shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
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
    print(f"attn_prefill bf16 B={B} H={H} T={T} causal={causal}: {us:8.1f} us  {fl / us * 1e-6:7.1f} TFLOP/s")
for (M, N, K, act, res) in ((16384, 2304, 768, 0, 0), (16384, 768, 768, 0, 1), (16384, 3072, 768, 2, 0), (16384, 768, 3072, 0, 1),
                            (32768, 3072, 768, 1, 0), (16384, 5120, 1280, 1, 0), (4096, 3072, 768, 2, 0)):
    A = torch.randn(M, K, device=dev, generator=g).bfloat16()
    W = (torch.randn(N, K, device=dev, generator=g) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev, generator=g)
    r = torch.randn(M, N, device=dev, generator=g) if res else None
    out = torch.empty(M, N, device=dev, dtype=torch.float32 if res else torch.bfloat16)
    us = timeit(lambda: check(lib.tutor_linear(1, A.data_ptr(), W.data_ptr(), bias.data_ptr(), r.data_ptr() if res else None,
                                               out.data_ptr(), 1 if res else 0, M, N, K, act, 0, None)))
    print(f"linear bf16 M={M} N={N} K={K} act={act} res={res}: {us:8.1f} us  {2.0 * M * N * K / us * 1e-6:7.1f} TFLOP/s")
