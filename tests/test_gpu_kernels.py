"""-m gpu: every CUDA kernel against a plain torch fp32 expression of the ATen op it stands in for
(SURVEY.md §2.2), called through the C ABI with device pointers."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
F32, BF16 = 0, 1


def ptr(t):
    return None if t is None else t.data_ptr()


def sync():
    torch.cuda.synchronize()


def rel_err(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-12)).item()


@pytest.mark.parametrize("M,d,eps", [(5, 128, 1e-5), (64, 768, 1e-5), (33, 1280, 1e-5), (16, 768, 1e-12),
                                     (2048, 768, 1e-12), (1501, 1280, 1e-5)])  # >= 1024 rows: the streaming flavour
def test_layernorm(pkg, lib, M, d, eps):
    g = torch.Generator(device="cpu").manual_seed(0)
    x = (torch.randn(M, d, generator=g) * 3 + 0.5).to(DEV)
    w, b = torch.randn(d, generator=g).to(DEV), torch.randn(d, generator=g).to(DEV)
    of, ob = torch.empty_like(x), torch.empty(M, d, dtype=torch.bfloat16, device=DEV)
    pkg._lib.check(lib.tutor_layernorm(ptr(x), ptr(w), ptr(b), eps, ptr(of), ptr(ob), M, d, None))
    sync()
    ref = torch.nn.functional.layer_norm(x, (d,), w, b, eps)
    assert (of - ref).abs().max().item() < 2e-5
    assert (ob.float() - ref).abs().max().item() < 0.05


def _linear_ref(A, W, bias, res, act):
    y = A.float() @ W.float().T
    if bias is not None:
        y = y + bias
    if act == 1:
        y = 0.5 * y * (1 + torch.tanh(math.sqrt(2 / math.pi) * (y + 0.044715 * y ** 3)))
    elif act == 2:
        y = torch.nn.functional.gelu(y)
    if res is not None:
        y = y + res
    return y


@pytest.mark.parametrize("M,N,K,act,use_res", [(1, 768, 768, 0, True), (3, 2304, 768, 0, False), (8, 3072, 768, 1, False),
                                               (1, 768, 3072, 0, True), (37, 200, 128, 2, True),
                                               (256, 384, 512, 1, False),
                                               # 9..64 rows: the register-blocked weight-streaming kernel (4 x 8 per warp)
                                               (32, 3072, 768, 1, False), (64, 768, 3072, 0, True), (9, 2304, 768, 0, False),
                                               (33, 50, 256, 2, True), (17, 1001, 128, 0, True)])
def test_linear_f32(pkg, lib, M, N, K, act, use_res):
    g = torch.Generator().manual_seed(M * 7 + N)
    A, W = torch.randn(M, K, generator=g).to(DEV), (torch.randn(N, K, generator=g) * 0.05).to(DEV)
    bias = torch.randn(N, generator=g).to(DEV)
    res = torch.randn(M, N, generator=g).to(DEV) if use_res else None
    out = torch.empty(M, N, device=DEV)
    pkg._lib.check(lib.tutor_linear(F32, ptr(A), ptr(W), ptr(bias), ptr(res), ptr(out), 1, M, N, K, act, 0, None))
    sync()
    ref = _linear_ref(A.double(), W.double(), bias.double(), None if res is None else res.double(), act).float()
    assert rel_err(out, ref) < 2e-6


# (M, N, K): decode-shaped (swap-AB, M <= 128) and prefill-shaped (row-major, M > 128), ragged edges included
BF16_SHAPES = [(1, 768, 768), (16, 2304, 768), (64, 768, 3072), (64, 3072, 768), (50, 200, 128), (128, 256, 64),
               (129, 256, 64), (300, 2304, 768), (1024, 3072, 768), (257, 100, 192), (300, 768, 3072), (512, 264, 2048)]


@pytest.mark.parametrize("impl", [1, 0], ids=["mma_sync_check", "tcgen05"])
@pytest.mark.parametrize("M,N,K", BF16_SHAPES)
def test_linear_bf16(pkg, lib, M, N, K, impl):
    g = torch.Generator().manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g).to(DEV).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(DEV).bfloat16()
    bias = torch.randn(N, generator=g).to(DEV)
    res = torch.randn(M, N, generator=g).to(DEV)
    guard = 8192  # guard zones around the output: the CTA-pair kernel's 256-row tiles overhang ragged M by up to 255 rows
    slab = torch.full((M * N + 2 * guard,), float("nan"), device=DEV)
    slab[:guard] = 7.0
    slab[guard + M * N:] = 7.0
    out = slab[guard:guard + M * N].view(M, N)
    pkg._lib.check(lib.tutor_linear(BF16, ptr(A), ptr(W), ptr(bias), ptr(res), ptr(out), 1, M, N, K, 1, impl, None))
    sync()
    assert (slab[:guard] == 7.0).all() and (slab[guard + M * N:] == 7.0).all(), "stores outside the output tensor"
    ref = _linear_ref(A.double(), W.double(), bias.double(), res.double(), 1).float()
    # bf16 inputs are exact in fp32, so only the accumulation order differs - except that the tcgen05 kernel's
    # row-major (M > 128) epilogue uses the one-instruction tanh (|err| ~ 5e-4), stated tolerance 2e-3
    assert rel_err(out, ref) < (2e-3 if impl == 0 and M > 128 else 1e-5)
    outb = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    pkg._lib.check(lib.tutor_linear(BF16, ptr(A), ptr(W), ptr(bias), None, ptr(outb), 0, M, N, K, 2, impl, None))
    sync()
    refb = _linear_ref(A.double(), W.double(), bias.double(), None, 2).float()
    assert rel_err(outb, refb) < 1e-2


@pytest.mark.parametrize("M,N,K", BF16_SHAPES)
def test_linear_bf16_stream_residual(pkg, lib, M, N, K):
    """The bf16 activation stream's linear layer: bf16(A W^T + bias + bf16 residual), every kernel flavour (swap-AB,
    single-CTA and CTA-pair TMA-staged epilogues, ragged edges), guard zones around the output."""
    g = torch.Generator().manual_seed(M + 3 * N + K)
    A = torch.randn(M, K, generator=g).to(DEV).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(DEV).bfloat16()
    bias = torch.randn(N, generator=g).to(DEV)
    res = (torch.randn(M, N, generator=g) * 2).to(DEV).bfloat16()
    guard = 8192
    slab = torch.full((M * N + 2 * guard,), float("nan"), device=DEV).bfloat16()
    slab[:guard] = 7.0
    slab[guard + M * N:] = 7.0
    out = slab[guard:guard + M * N].view(M, N)
    pkg._lib.check(lib.tutor_linear_bf16_residual(ptr(A), ptr(W), ptr(bias), ptr(res), ptr(out), M, N, K, 0, None))
    sync()
    assert (slab[:guard] == 7.0).all() and (slab[guard + M * N:] == 7.0).all(), "stores outside the output tensor"
    ref = _linear_ref(A.double(), W.double(), bias.double(), res.double(), 0)
    # one bf16 rounding of the fp32 sum: half an ulp = 2^-9 relative, plus the accumulation-order difference
    assert ((out.double() - ref).abs() <= ref.abs() * 2.0 ** -8 + 1e-4).all()


@pytest.mark.parametrize("M,d,eps", [(5, 128, 1e-12), (64, 768, 1e-12), (1025, 768, 1e-12), (33, 1280, 1e-5), (7, 2048, 1e-5)])
def test_layernorm_bf16_rows(pkg, lib, M, d, eps):
    g = torch.Generator(device="cpu").manual_seed(1)
    x = (torch.randn(M, d, generator=g) * 3 + 0.5).to(DEV).bfloat16()
    w, b = torch.randn(d, generator=g).to(DEV), torch.randn(d, generator=g).to(DEV)
    of, ob = torch.empty(M, d, device=DEV), torch.empty(M, d, dtype=torch.bfloat16, device=DEV)
    pkg._lib.check(lib.tutor_layernorm_bf16(ptr(x), ptr(w), ptr(b), eps, ptr(of), ptr(ob), M, d, None))
    sync()
    ref = torch.nn.functional.layer_norm(x.float(), (d,), w, b, eps)
    assert (of - ref).abs().max().item() < 2e-5
    assert (ob.float() - ref).abs().max().item() < 0.05
    ob2 = torch.empty_like(ob)
    pkg._lib.check(lib.tutor_layernorm_bf16(ptr(x), ptr(w), ptr(b), eps, None, ptr(ob2), M, d, None))
    sync()
    assert torch.equal(ob, ob2)


def _sdpa_ref(q, k, v, causal):
    # q [B,T,H,64], k/v [B,H,S,64]
    qh = q.permute(0, 2, 1, 3).double()
    s = qh @ k.double().transpose(-1, -2) / 8.0
    if causal:
        T, S = s.shape[-2:]
        mask = torch.ones(T, S, dtype=torch.bool, device=s.device).tril(S - T)
        s = s.masked_fill(~mask, float("-inf"))
    o = torch.softmax(s, -1) @ v.double()
    return o.permute(0, 2, 1, 3).float()


@pytest.mark.parametrize("dtype", [F32, BF16])
@pytest.mark.parametrize("B,H,T,causal", [(1, 2, 32, 1), (2, 12, 100, 1), (3, 2, 40, 0), (2, 4, 129, 1), (1, 12, 512, 0),
                                          (1, 1, 128, 0), (2, 3, 300, 0), (2, 2, 257, 1), (1, 20, 1024, 1), (1, 2, 1, 0)])
def test_attn_prefill(pkg, lib, dtype, B, H, T, causal):
    g = torch.Generator().manual_seed(T)
    td = torch.float32 if dtype == F32 else torch.bfloat16
    Smax = T + 7
    q = torch.randn(B, T, H, 64, generator=g).to(DEV).to(td)
    k = torch.randn(B, H, Smax, 64, generator=g).to(DEV).to(td)
    v = torch.randn(B, H, Smax, 64, generator=g).to(DEV).to(td)
    # the output sits between two guard zones: a kernel that stores past either end of it (query tiles are 128 rows, the
    # last one is ragged) would change them
    n, guard = B * T * H * 64, 4096
    slab = torch.full((n + 2 * guard,), 7.0, device=DEV, dtype=td)
    out = slab[guard:guard + n].view(B, T, H, 64)
    kv_before = (k.clone(), v.clone())
    pkg._lib.check(lib.tutor_attn_prefill(dtype, ptr(q), ptr(k), ptr(v), ptr(out), B, H, T, Smax, causal, None))
    sync()
    ref = _sdpa_ref(q, k[:, :, :T], v[:, :, :T], causal)
    tol = 2e-5 if dtype == F32 else 2e-2
    assert (out.float() - ref).abs().max().item() < tol
    assert (slab[:guard] == 7.0).all() and (slab[guard + n:] == 7.0).all(), "stores outside the output tensor"
    assert torch.equal(k, kv_before[0]) and torch.equal(v, kv_before[1]), "the cache must not be written"


@pytest.mark.parametrize("dtype", [F32, BF16])
@pytest.mark.parametrize("B,H,S", [(1, 12, 1), (1, 12, 33), (4, 2, 95), (64, 12, 513), (3, 20, 1024)])
def test_attn_decode(pkg, lib, dtype, B, H, S):
    g = torch.Generator().manual_seed(S)
    td = torch.float32 if dtype == F32 else torch.bfloat16
    Smax = (S + 63) // 64 * 64
    q = torch.randn(B, 1, H, 64, generator=g).to(DEV).to(td)
    k = torch.randn(B, H, Smax, 64, generator=g).to(DEV).to(td)
    v = torch.randn(B, H, Smax, 64, generator=g).to(DEV).to(td)
    out = torch.empty(B, 1, H, 64, device=DEV, dtype=td)
    pkg._lib.check(lib.tutor_attn_decode(dtype, ptr(q), ptr(k), ptr(v), ptr(out), B, H, Smax, S, None))
    sync()
    ref = _sdpa_ref(q, k[:, :, :S], v[:, :, :S], 0)
    tol = 2e-5 if dtype == F32 else 2e-2
    assert (out.float() - ref).abs().max().item() < tol


# B*H = 768 / 320 / 520: the two-, four- and three-stage rings of the TMA kernel
@pytest.mark.parametrize("B,H,S,ks", [(1, 2, 5, 1), (5, 12, 130, 3), (64, 12, 640, 6), (16, 20, 961, 4), (26, 20, 417, 2),
                                      (20, 12, 64, 2)])
def test_attn_decode_fused(pkg, lib, B, H, S, ks):
    """The decode step's own attention instance: split-K partials of c_attn summed in order + bias, K/V row appended at
    position S-1, attention over S rows - against fp64 SDPA on the same (bf16-rounded) q, k, v."""
    g = torch.Generator().manual_seed(S + ks)
    d = H * 64
    Smax = (S + 63) // 64 * 64
    part = (torch.randn(ks, B, 3 * d, generator=g) * 0.5).to(DEV)
    bias = (torch.randn(3 * d, generator=g) * 0.1).to(DEV)
    k = torch.randn(B, H, Smax, 64, generator=g).to(DEV).bfloat16()
    v = torch.randn(B, H, Smax, 64, generator=g).to(DEV).bfloat16()
    k0, v0 = k.clone(), v.clone()
    out = torch.empty(B, d, device=DEV, dtype=torch.bfloat16)
    pkg._lib.check(lib.tutor_attn_decode_fused(ptr(part), ks, ptr(bias), ptr(k), ptr(v), ptr(out), B, H, Smax, S, None))
    sync()
    qkv = bias.clone()[None].repeat(B, 1)
    for z in range(ks):  # the kernel's summation order
        qkv = qkv + part[z]
    q, kn, vn = qkv[:, :d], qkv[:, d:2 * d].bfloat16(), qkv[:, 2 * d:].bfloat16()
    assert torch.equal(k[:, :, S - 1], kn.view(B, H, 64)) and torch.equal(v[:, :, S - 1], vn.view(B, H, 64)), "cache append"
    assert torch.equal(k[:, :, :S - 1], k0[:, :, :S - 1]) and torch.equal(k[:, :, S:], k0[:, :, S:]), "other rows untouched"
    ref = _sdpa_ref(q.view(B, 1, H, 64), k[:, :, :S], v[:, :, :S], 0)
    assert (out.float().view(B, 1, H, 64) - ref).abs().max().item() < 2e-2


@pytest.mark.parametrize("dtype,B,V,d", [(F32, 1, 50257, 768), (F32, 4, 1000, 128), (F32, 11, 5000, 256),
                                         (BF16, 1, 1000, 128), (BF16, 64, 50257, 768), (BF16, 16, 50257, 1280),
                                         (BF16, 100, 3000, 128)])
@pytest.mark.parametrize("penalty", [1.0, 1.2])
def test_lmhead_argmax(pkg, lib, dtype, B, V, d, penalty):
    g = torch.Generator().manual_seed(V + B)
    td = torch.float32 if dtype == F32 else torch.bfloat16
    x = torch.randn(B, d, generator=g).to(DEV).to(td)
    wte = (torch.randn(V, d, generator=g) * 0.02).to(DEV).to(td)
    logits = x.double() @ wte.double().T
    words = (V + 31) // 32
    seen_ids = torch.topk(logits, 3, dim=1).indices  # penalise exactly the would-be winners
    seen = torch.zeros(B, words, dtype=torch.int32, device=DEV)
    for b in range(B):
        for t in seen_ids[b].tolist():
            seen[b, t // 32] |= (1 << (t % 32)) if (t % 32) < 31 else -(1 << 31)
    if penalty != 1.0:
        mask = torch.zeros(B, V, dtype=torch.bool, device=DEV).scatter_(1, seen_ids, True)
        pen = torch.where(logits < 0, logits * penalty, logits / penalty)
        logits = torch.where(mask, pen, logits)
    ref = logits.argmax(-1).to(torch.int32)
    out = torch.empty(B, dtype=torch.int32, device=DEV)
    scratch = torch.empty(lib.tutor_lmhead_scratch_bytes(B), dtype=torch.uint8, device=DEV)
    pkg._lib.check(lib.tutor_lmhead_argmax(dtype, ptr(x), ptr(wte), B, V, d, ptr(seen), penalty, ptr(out), ptr(scratch),
                                           None))
    sync()
    bad = (out != ref).nonzero().flatten().tolist()
    for b in bad:  # a mismatch is only acceptable on a numerical near-tie
        assert abs(logits[b, out[b]].item() - logits[b, ref[b]].item()) < 1e-4 * logits[b].abs().max().item(), b


def test_lmhead_tie_breaks_to_lowest_id(pkg, lib):
    B, V, d = 2, 1000, 128
    x = torch.zeros(B, d, device=DEV)
    x[:, 0] = 1.0
    wte = torch.zeros(V, d, device=DEV)
    wte[[7, 500, 999], 0] = 3.0  # three exact ties
    out = torch.empty(B, dtype=torch.int32, device=DEV)
    scratch = torch.empty(lib.tutor_lmhead_scratch_bytes(B), dtype=torch.uint8, device=DEV)
    for dtype, xx, ww in ((F32, x, wte), (BF16, x.bfloat16(), wte.bfloat16())):
        pkg._lib.check(lib.tutor_lmhead_argmax(dtype, ptr(xx), ptr(ww), B, V, d, None, 1.0, ptr(out), ptr(scratch), None))
        sync()
        assert out.tolist() == [7, 7]  # torch.argmax returns the first maximal index


def test_meanpool_and_cosine(pkg, lib):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(3, 77, 768, generator=g).to(DEV)
    out = torch.empty(3, 768, device=DEV)
    pkg._lib.check(lib.tutor_meanpool(ptr(x), ptr(out), 3, 77, 768, None))
    sync()
    assert (out - x.mean(1)).abs().max().item() < 1e-6
    a, b = torch.randn(9, 768, generator=g).to(DEV), torch.randn(9, 768, generator=g).to(DEV)
    a[4] = 0  # eps path: max(|x|, 1e-8)
    cs = torch.empty(9, device=DEV)
    pkg._lib.check(lib.tutor_cosine_pairs(ptr(a), ptr(b), ptr(cs), 9, 768, None))
    sync()
    ref = torch.nn.functional.cosine_similarity(a, b)
    assert (cs - ref).abs().max().item() < 1e-6
    assert abs(pkg.cosine_similarity(a[:1].cpu(), b[:1].cpu()) - ref[0].item()) < 1e-6


# ---- decode-step linears with fused LayerNorm operand / cluster split-K reduction (gemm_decode_fused.cu)
def _tile_stats(lib, pkg, x):
    M, d = x.shape
    stats = torch.empty(d // 128, M, 2, device=DEV)
    pkg._lib.check(lib.tutor_row_tile_stats(ptr(x), ptr(stats), M, d, None))
    return stats


def _ln_bf16(x, g, b, eps):
    # what the kernel feeds the tensor core: LayerNorm in fp32, rounded to bf16
    return torch.nn.functional.layer_norm(x, (x.shape[1],), g, b, eps).bfloat16()


def test_row_tile_stats(pkg, lib):
    g = torch.Generator().manual_seed(11)
    x = (torch.randn(37, 768, generator=g) * 2 + 0.3).to(DEV)
    stats = _tile_stats(lib, pkg, x)
    sync()
    xt = x.view(37, 6, 128).permute(1, 0, 2).double()
    assert (stats[..., 0] - xt.mean(-1)).abs().max().item() < 1e-6
    assert rel_err(stats[..., 1], ((xt - xt.mean(-1, keepdim=True)) ** 2).sum(-1)) < 1e-5


# (M, N, K, ks, bn): engine shapes for GPT-2 small / large, ragged batches, every batch tile, ring wrap (nk/ks > stages)
LN_ACT_SHAPES = [(64, 3072, 768, 6, 64), (64, 3072, 768, 0, 0), (16, 5120, 1280, 2, 16), (50, 3072, 768, 3, 32),
                 (1, 256, 128, 2, 16), (100, 384, 768, 1, 128), (33, 128, 1280, 4, 64), (128, 1024, 512, 8, 128)]


@pytest.mark.parametrize("M,N,K,ks,bn", LN_ACT_SHAPES)
def test_decode_ln_linear_act(pkg, lib, M, N, K, ks, bn):
    g = torch.Generator().manual_seed(M + N + K)
    x = (torch.randn(M, K, generator=g) * 1.7 + 0.2).to(DEV)
    gam, bet = (1 + 0.1 * torch.randn(K, generator=g)).to(DEV), (0.1 * torch.randn(K, generator=g)).to(DEV)
    W = (torch.randn(N, K, generator=g) * 0.05).to(DEV).bfloat16()
    bias = torch.randn(N, generator=g).to(DEV)
    stats = _tile_stats(lib, pkg, x)
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    pkg._lib.check(lib.tutor_decode_ln_linear_act(ptr(x), ptr(stats), ptr(gam), ptr(bet), 1e-5, ptr(W), ptr(bias), 1, M, N,
                                                  K, ks, bn, ptr(out), None))
    sync()
    ref = _linear_ref(_ln_bf16(x, gam, bet, 1e-5).double(), W.double(), bias.double(), None, 1).float()
    # bf16 rounding of LN(x) can differ by one ulp from torch's for values on a rounding boundary; output is bf16
    assert rel_err(out, ref) < 1e-2


@pytest.mark.parametrize("M,N,K,ks", [(64, 2304, 768, 6), (64, 2304, 768, 0), (16, 3840, 1280, 4), (7, 384, 256, 2),
                                      (128, 256, 768, 12)])
def test_decode_ln_linear_partial(pkg, lib, M, N, K, ks):
    g = torch.Generator().manual_seed(M + N + K + 1)
    x = (torch.randn(M, K, generator=g) * 1.3 - 0.1).to(DEV)
    gam, bet = (1 + 0.1 * torch.randn(K, generator=g)).to(DEV), (0.1 * torch.randn(K, generator=g)).to(DEV)
    W = (torch.randn(N, K, generator=g) * 0.05).to(DEV).bfloat16()
    stats = _tile_stats(lib, pkg, x)
    nsplit = ks if ks else 12
    partial = torch.zeros(nsplit, M, N, device=DEV)
    pkg._lib.check(lib.tutor_decode_ln_linear_partial(ptr(x), ptr(stats), ptr(gam), ptr(bet), 1e-5, ptr(W), M, N, K, ks,
                                                      ptr(partial), None))
    sync()
    ref = _ln_bf16(x, gam, bet, 1e-5).double() @ W.double().T
    assert rel_err(partial.double().sum(0), ref) < 2e-3  # a one-ulp bf16 difference in LN(x) shows at ~1e-3


@pytest.mark.parametrize("M,N,K,ks,bn", [(64, 768, 768, 6, 32), (64, 768, 3072, 8, 32), (64, 768, 3072, 0, 0),
                                         (64, 768, 768, 6, 16), (16, 1280, 5120, 8, 16), (50, 256, 512, 4, 64),
                                         (1, 128, 128, 2, 16), (128, 384, 1024, 8, 128), (100, 768, 768, 3, 32)])
def test_decode_linear_residual(pkg, lib, M, N, K, ks, bn):
    g = torch.Generator().manual_seed(M + N + K + 2)
    A = torch.randn(M, K, generator=g).to(DEV).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(DEV).bfloat16()
    bias = torch.randn(N, generator=g).to(DEV)
    x0 = torch.randn(M, N, generator=g).to(DEV)
    x = x0.clone()
    stats = torch.full((N // 128, M, 2), float("nan"), device=DEV)
    pkg._lib.check(lib.tutor_decode_linear_residual(ptr(A), ptr(W), ptr(bias), M, N, K, ks, bn, ptr(x), ptr(stats), None))
    sync()
    ref = _linear_ref(A.double(), W.double(), bias.double(), x0.double(), 0)
    assert rel_err(x, ref.float()) < 1e-5
    rt = ref.view(M, N // 128, 128).permute(1, 0, 2)
    assert (stats[..., 0] - rt.mean(-1)).abs().max().item() < 1e-5
    assert rel_err(stats[..., 1], ((rt - rt.mean(-1, keepdim=True)) ** 2).sum(-1)) < 1e-4
    # deterministic: a second run from the same input gives the same bits
    x2 = x0.clone()
    pkg._lib.check(lib.tutor_decode_linear_residual(ptr(A), ptr(W), ptr(bias), M, N, K, ks, bn, ptr(x2), ptr(stats), None))
    sync()
    assert torch.equal(x, x2)


def test_decode_fused_chain_matches_layernorm(pkg, lib):
    """residual GEMM -> statistics -> LN-operand GEMM equals the unfused LayerNorm + linear on the same x."""
    g = torch.Generator().manual_seed(99)
    M, d = 64, 768
    A = torch.randn(M, d, generator=g).to(DEV).bfloat16()
    Wp = (torch.randn(d, d, generator=g) * 0.05).to(DEV).bfloat16()
    Wf = (torch.randn(4 * d, d, generator=g) * 0.05).to(DEV).bfloat16()
    bp, bf = torch.randn(d, generator=g).to(DEV), torch.randn(4 * d, generator=g).to(DEV)
    gam, bet = (1 + 0.1 * torch.randn(d, generator=g)).to(DEV), (0.1 * torch.randn(d, generator=g)).to(DEV)
    x = torch.randn(M, d, generator=g).to(DEV)
    stats = torch.empty(d // 128, M, 2, device=DEV)
    h = torch.empty(M, 4 * d, dtype=torch.bfloat16, device=DEV)
    pkg._lib.check(lib.tutor_decode_linear_residual(ptr(A), ptr(Wp), ptr(bp), M, d, d, 0, 0, ptr(x), ptr(stats), None))
    pkg._lib.check(lib.tutor_decode_ln_linear_act(ptr(x), ptr(stats), ptr(gam), ptr(bet), 1e-5, ptr(Wf), ptr(bf), 1, M,
                                                  4 * d, d, 0, 0, ptr(h), None))
    sync()
    ref = _linear_ref(_ln_bf16(x, gam, bet, 1e-5).double(), Wf.double(), bf.double(), None, 1).float()
    assert rel_err(h, ref) < 1e-2
