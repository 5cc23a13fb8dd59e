"""CPU, world_size 2 over gloo: the one exchange step of the sharded corpus search (configs[4]) — shard
bounds, the all-gather layout, and the merge rule — against an unsharded top-k."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def host_merge(scores, idx, k):
    """(score desc, index asc) top-k of [Q, n] candidates; idx < 0 marks an empty slot."""
    s = torch.where(idx < 0, torch.full_like(scores, float("-inf")), scores)
    order = torch.argsort(idx, dim=1, stable=True)
    s, i = torch.gather(s, 1, order), torch.gather(idx, 1, order)
    top = torch.argsort(s, dim=1, descending=True, stable=True)[:, :k]
    return torch.gather(s, 1, top), torch.gather(i, 1, top)


def _worker(rank, ws, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    import lms_b200 as pkg
    g = torch.Generator().manual_seed(0)
    corpus = torch.nn.functional.normalize(torch.randn(1001, 64, generator=g), dim=1)
    corpus[700] = corpus[3]  # an exact tie across the two shards: the lower row index must win
    q = torch.randn(5, 64, generator=g)
    k = 4
    lo, hi = pkg.shard_bounds(corpus.shape[0], ws, rank)
    local = torch.nn.functional.cosine_similarity(q[:, None], corpus[None, lo:hi], dim=2)
    ls, li = local.topk(k, dim=1)
    cand_s, cand_i = pkg.exchange_candidates(ls, (li + lo).to(torch.int32))
    ms, mi = host_merge(cand_s, cand_i, k)
    full = torch.nn.functional.cosine_similarity(q[:, None], corpus[None], dim=2)
    fs, fi = full.topk(k, dim=1)
    ok = torch.equal(mi.long(), fi) and torch.allclose(ms, fs)
    lo0, hi0 = pkg.shard_bounds(corpus.shape[0], ws, 0)
    ok = ok and tuple(cand_s.shape) == (5, ws * k) and bool((cand_i[:, :k] < hi0).all())  # rank-0 block first
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_exchange_and_merge_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    [p.start() for p in procs]
    [p.join(180) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert ret.get(0) and ret.get(1)
