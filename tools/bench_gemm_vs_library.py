"""The BERT-base / GPT-2 layer GEMMs at prefill size: this library's tcgen05 kernels (bias and activation in the epilogue,
bf16 out) against torch.matmul on the same operands (cuBLAS, no bias, no activation) - a yardstick, not a code path.
Each shape runs 20 times back to back inside CUDA events after 3 warm-up calls; operands of all shapes exceed L2 together."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200

lib, check = lms_b200._lib.load(), lms_b200._lib.check
M = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
SHAPES = [("c_attn / qkv", 2304, 768, 0), ("attention out", 768, 768, 0), ("intermediate + gelu", 3072, 768, 2),
          ("output", 768, 3072, 0), ("gpt2-large c_fc", 5120, 1280, 1), ("gpt2-large c_proj", 1280, 5120, 0)]
g = torch.Generator(device="cuda").manual_seed(0)


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


for name, N, K, act in SHAPES:
    A = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    W = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    Wt = W.t()
    st = torch.cuda.current_stream().cuda_stream
    mine = timed(lambda: check(lib.tutor_linear(1, A.data_ptr(), W.data_ptr(), bias.data_ptr(), None, out.data_ptr(), 0, M, N, K,
                                                act, 0, st)))
    mine_res = timed(lambda: check(lib.tutor_linear_bf16_residual(A.data_ptr(), W.data_ptr(), bias.data_ptr(), res.data_ptr(),
                                                                  out.data_ptr(), M, N, K, act, st)))
    libt = timed(lambda: torch.matmul(A, Wt, out=out))
    tf = 2.0 * M * N * K / 1e6
    print(f"{name:22s} [{M} x {N} x {K}]  this library {mine:7.1f} us ({tf / mine:6.0f} TFLOP/s)   + bf16 residual {mine_res:7.1f} us"
          f"   torch.matmul {libt:7.1f} us ({tf / libt:6.0f} TFLOP/s)", flush=True)
