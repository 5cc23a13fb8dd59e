"""Per-launch time of the fused decode linears (gemm_decode_fused.cu) for GPT-2 small shapes, each replayed
back to back inside a CUDA graph (no PDL, so this is launch + in-kernel time of a dependent chain)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200

lib = lms_b200._lib.load()
chk = lms_b200._lib.check
DEV = "cuda:0"
M, d = 64, 768
g = torch.Generator().manual_seed(0)
x = torch.randn(M, d, generator=g).to(DEV)
gam, bet = torch.ones(d, device=DEV), torch.zeros(d, device=DEV)
stats = torch.empty(d // 128, M, 2, device=DEV)
chk(lib.tutor_row_tile_stats(x.data_ptr(), stats.data_ptr(), M, d, None))
W = {n: (torch.randn(N, K, generator=g) * 0.02).to(DEV).bfloat16() for n, (N, K) in
     {"qkv": (3 * d, d), "proj": (d, d), "fc": (4 * d, d), "proj2": (d, 4 * d)}.items()}
bias = torch.zeros(4 * d, device=DEV)
att = torch.randn(M, d, generator=g).to(DEV).bfloat16()
hb = torch.randn(M, 4 * d, generator=g).to(DEV).bfloat16()
out_h = torch.empty(M, 4 * d, dtype=torch.bfloat16, device=DEV)
partial = torch.empty(12, M, 3 * d, device=DEV)
REP = 40


def timed(name, fn):
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn(st.cuda_stream)
        st.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for _ in range(REP):
                fn(st.cuda_stream)
        for _ in range(3):
            gr.replay()
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(10):
            gr.replay()
        e1.record(st)
        st.synchronize()
    print(f"{name:48s} {e0.elapsed_time(e1) * 1000 / (10 * REP):7.2f} us/launch", flush=True)


for ks in (6, 12):
    timed(f"c_attn LN+partial ks={ks}", lambda s, ks=ks: chk(lib.tutor_decode_ln_linear_partial(
        x.data_ptr(), stats.data_ptr(), gam.data_ptr(), bet.data_ptr(), 1e-5, W["qkv"].data_ptr(), M, 3 * d, d, ks,
        partial.data_ptr(), s)))
for ks, bn in ((6, 64), (3, 64), (1, 64), (6, 32), (3, 32), (2, 16)):
    timed(f"c_fc LN+cluster gelu ks={ks} bn={bn}", lambda s, ks=ks, bn=bn: chk(lib.tutor_decode_ln_linear_act(
        x.data_ptr(), stats.data_ptr(), gam.data_ptr(), bet.data_ptr(), 1e-5, W["fc"].data_ptr(), bias.data_ptr(), 1, M,
        4 * d, d, ks, bn, out_h.data_ptr(), s)))
for ks, bn in ((6, 64), (6, 32), (6, 16), (3, 16), (2, 16), (1, 16), (1, 64)):
    timed(f"proj cluster resid ks={ks} bn={bn}", lambda s, ks=ks, bn=bn: chk(lib.tutor_decode_linear_residual(
        att.data_ptr(), W["proj"].data_ptr(), bias.data_ptr(), M, d, d, ks, bn, x.data_ptr(), stats.data_ptr(), s)))
for ks, bn in ((8, 64), (8, 32), (8, 16), (6, 16), (4, 16), (4, 32), (2, 16)):
    timed(f"proj2 cluster resid ks={ks} bn={bn}", lambda s, ks=ks, bn=bn: chk(lib.tutor_decode_linear_residual(
        hb.data_ptr(), W["proj2"].data_ptr(), bias.data_ptr(), M, d, 4 * d, ks, bn, x.data_ptr(), stats.data_ptr(), s)))
