"""Decode attention at configs[2] shape (or B H S Smax from the command line, e.g. 16 20 960 1024 for configs[3]): the fused instance (c_attn partials reduced in the prologue, K/V row appended) at
ks = 2 and 6 against the plain instance (q given), each as 12 launches over 12 layers of cache in one CUDA graph."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
lib, check = lms_b200._lib.load(), lms_b200._lib.check
B, H, L, S, Smax = 64, 12, 12, 640, 768
if len(sys.argv) >= 5:
    B, H, S, Smax = (int(a) for a in sys.argv[1:5])
if len(sys.argv) == 6:
    L = int(sys.argv[5])  # 1: every launch streams the same layer (L2-resident when it fits)
d = H * 64
g = torch.Generator(device="cuda").manual_seed(0)
k = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
v = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
q = torch.randn(B, d, device="cuda", generator=g).bfloat16()
bias = torch.randn(3 * d, device="cuda", generator=g) * 0.1
out = torch.empty(B, d, device="cuda", dtype=torch.bfloat16)

def timed(launch, name):
    side = torch.cuda.Stream()
    for l in range(L):
        launch(l, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            for l in range(L):
                launch(l, side.cuda_stream)
        graph.replay(); side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(10):
            graph.replay()
        e1.record(side); side.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (10 * L)
    kv = B * H * (S - 1) * 64 * 2 * 2
    print(f"{name:44s} {us:6.2f} us per launch, {kv / us / 1e3:6.0f} GB/s of K/V rows ({kv / us / 1e3 / 6549.1:.3f} of the HBM peak)", flush=True)

for ks in (2, 6):
    part = torch.randn(ks, B, 3 * d, device="cuda", generator=g) * 0.3
    timed(lambda l, st: check(lib.tutor_attn_decode_fused(part.data_ptr(), ks, bias.data_ptr(), k[l].data_ptr(), v[l].data_ptr(),
                                                          out.data_ptr(), B, H, Smax, S, st)), f"fused instance, ks = {ks}")
timed(lambda l, st: check(lib.tutor_attn_decode(1, q.data_ptr(), k[l].data_ptr(), v[l].data_ptr(), out.data_ptr(), B, H, Smax, S, st)),
      "plain instance (q given, S rows cached)")
