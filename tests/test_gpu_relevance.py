"""-m gpu: corpus cosine top-k (configs[4]) against the CPU oracle; full-size run checked through
size-independent properties."""
import pytest
import torch

from oracle import reference_path as rp

pytestmark = pytest.mark.gpu


def corpus(n, d=768, seed=3):
    g = torch.Generator().manual_seed(seed)
    c = torch.randn(n, d, generator=g)
    return c / c.norm(dim=1, keepdim=True)


# the last two leave the tensor-core path: more than 32 queries (SIMT scan, register-row kernel) and k > 32 (SIMT scan,
# block-wide chunk selection)
@pytest.mark.parametrize("N,Q,k,d", [(50_000, 32, 10, 768), (250_000, 32, 10, 768), (4097, 3, 5, 768), (7, 2, 10, 128), (20_000, 1, 1, 128),
                                     (3000, 40, 5, 768), (5000, 4, 40, 128), (9000, 32, 32, 768)])
def test_topk_matches_oracle(pkg, N, Q, k, d):
    c = corpus(N, d)
    q = torch.randn(Q, d, generator=torch.Generator().manual_seed(9)) * 3
    shard = pkg.CorpusShard(c.cuda(), row_lo=0)
    s, i = shard.topk(q.cuda(), k)
    torch.cuda.synchronize()
    rs, ri = rp.corpus_topk(q, c, min(k, N))
    kk = min(k, N)
    assert torch.equal(i[:, :kk].cpu().long(), ri)
    assert (s[:, :kk].cpu() - rs).abs().max().item() < 1e-4
    if k > N:
        assert (i[:, N:] == -1).all()


def test_sharded_equals_unsharded_single_gpu(pkg):
    """Emulates the 8-rank layout on one GPU: per-shard top-k, candidates concatenated in rank order, device merge."""
    N, Q, k = 100_003, 8, 10
    c, q = corpus(N).cuda(), torch.randn(Q, 768, generator=torch.Generator().manual_seed(1)).cuda()
    full_s, full_i = pkg.CorpusShard(c).topk(q, k)
    cs, ci = [], []
    for r in range(8):
        lo, hi = pkg.shard_bounds(N, 8, r)
        s, i = pkg.CorpusShard(c[lo:hi], row_lo=lo).topk(q, k)
        cs.append(s); ci.append(i)
    cand_s, cand_i = torch.cat(cs, 1).contiguous(), torch.cat(ci, 1).contiguous()
    out_s, out_i = torch.empty_like(full_s), torch.empty_like(full_i)
    lib = pkg._lib.load()
    pkg._lib.check(lib.tutor_topk_merge(cand_s.data_ptr(), cand_i.data_ptr(), Q, 80, k, out_s.data_ptr(), out_i.data_ptr(),
                                        None))
    torch.cuda.synchronize()
    assert torch.equal(out_i, full_i) and torch.equal(out_s, full_s)


def test_near_ties_around_rank_k_fall_back_to_the_exact_scan(pkg):
    """The plain-TF32 filter scan cannot rank rows whose cosines differ by less than its error (~2e-3): 200 planted rows
    spaced 1e-4 apart around the top of query 0's ranking, each in a different 32-row group.  The library must notice that
    its candidate set is not provably complete and run the error-compensated scan: index sets identical to the oracle."""
    N, Q, k, d = 220_000, 4, 10, 768  # the filter scan is used from 200 k rows on
    g = torch.Generator().manual_seed(11)
    c = corpus(N, d, seed=12)
    q = torch.randn(Q, d, generator=g)
    q0 = q[0] / q[0].norm()
    for j in range(200):
        u = torch.randn(d, generator=g)
        u = u - (u @ q0) * q0
        u = u / u.norm()
        cos = 0.9 - 1e-4 * j
        c[j * 577 + 5] = cos * q0 + (1 - cos * cos) ** 0.5 * u
    s, i = pkg.CorpusShard(c.cuda()).topk(q.cuda(), k)
    torch.cuda.synchronize()
    rs, ri = rp.corpus_topk(q, c, k)
    assert torch.equal(i.cpu().long(), ri)
    assert (s.cpu() - rs).abs().max().item() < 1e-4
    assert ri[0].tolist() == [j * 577 + 5 for j in range(k)]


def test_full_size_properties(pkg):
    """1M x 768 fp32 corpus (3.07 GB): planted rows must be found, scores sorted, self-similarity 1."""
    N, Q, k = 1_000_000, 32, 10
    g = torch.Generator(device="cuda").manual_seed(3)
    c = torch.randn(N, 768, generator=g, device="cuda")
    q = torch.randn(Q, 768, generator=g, device="cuda")
    planted = torch.arange(Q, device="cuda") * 31_111 + 17
    c[planted] = q * 2.5  # cosine is scale-invariant: the planted row scores exactly ~1
    s, i = pkg.CorpusShard(c).topk(q, k)
    torch.cuda.synchronize()
    assert torch.equal(i[:, 0].long(), planted)
    assert (s[:, 0] - 1).abs().max().item() < 1e-5
    assert (s[:, :-1] >= s[:, 1:]).all()
    assert (i >= 0).all() and (i < N).all()


def test_tensor_core_scan_is_fp32_accurate(pkg):
    """The 3xTF32 tensor-core scan must be as accurate as an fp32 FMA chain: top-k scores against float64."""
    N, Q, k = 30_000, 32, 10
    c = corpus(N) * torch.rand(N, 1, generator=torch.Generator().manual_seed(4)) * 7  # un-normalised rows
    q = torch.randn(Q, 768, generator=torch.Generator().manual_seed(5))
    s, i = pkg.CorpusShard(c.cuda()).topk(q.cuda(), k)
    torch.cuda.synchronize()
    ref = torch.nn.functional.cosine_similarity(q.double()[:, None, :], c.double()[None, :, :], dim=2)
    rs, ri = ref.topk(k, dim=1)
    assert torch.equal(i.cpu().long(), ri)
    assert (s.cpu().double() - rs).abs().max().item() < 2e-6
