"""torchrun check of the sharded corpus search (configs[4]): every rank scans its row shard, one NCCL all-gather of the
[Q,k] candidates, device merge; rank 0 compares with the unsharded oracle top-k.  Also times the search."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import lms_b200
from oracle import reference_path as rp

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N, Q, k = int(os.environ.get("CORPUS_ROWS", 1_000_000)), 32, 10
g = torch.Generator().manual_seed(3)
lo, hi = lms_b200.shard_bounds(N, world, rank)
# every rank draws the same stream and keeps its own slice (small enough on the host in 100k-row blocks)
rows = []
for start in range(0, N, 100_000):
    blk = torch.randn(min(100_000, N - start), 768, generator=g)
    a, b = max(lo, start), min(hi, start + blk.shape[0])
    if a < b:
        rows.append(blk[a - start:b - start])
queries = torch.randn(Q, 768, generator=g)
shard = lms_b200.CorpusShard(torch.cat(rows).cuda(), row_lo=lo)
qd = queries.cuda()
s_py, i_py = shard.search(qd, k)        # exchange through torch.distributed (two all-gathers + layout copies)
lms_b200.init_exchange()                 # from here on: the library's own ncclAllGather of packed candidates
s, i = shard.search(qd, k)
torch.cuda.synchronize()
assert torch.equal(i, i_py) and torch.equal(s, s_py), "the two exchange paths disagree"
dist.barrier()


def timed(fn, n=20):
    fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t


ms = timed(lambda: shard.search(qd, k))
# the same search captured once and replayed: scan -> top-k -> pack -> all-gather -> merge as one CUDA graph
side = torch.cuda.Stream()
graph = torch.cuda.CUDAGraph()
with torch.cuda.stream(side):
    shard.search(qd, k)
    side.synchronize()
    with torch.cuda.graph(graph, stream=side):
        gs, gi = shard.search(qd, k)
    ms_graph = timed(graph.replay)
assert torch.equal(gi, i) and torch.equal(gs, s)
if rank == 0:
    print(f"eager {ms.item() * 1e3:.1f} us, CUDA graph {ms_graph.item() * 1e3:.1f} us per search")
ms = ms_graph
if rank == 0:
    g = torch.Generator().manual_seed(3)
    full = torch.cat([torch.randn(min(100_000, N - st), 768, generator=g) for st in range(0, N, 100_000)])
    rs, ri = rp.corpus_topk(queries, full, k)
    ok = torch.equal(i.cpu().long(), ri) and (s.cpu() - rs).abs().max().item() < 1e-4
    print(f"sharded search over {world} GPUs: {'OK' if ok else 'MISMATCH'}; {ms.item() * 1e3:.1f} us per 32-query search "
          f"(max over ranks), {N * 768 * 4 / world / (ms.item() * 1e-3) / 1e9:.0f} GB/s per GPU")
    assert ok
dist.destroy_process_group()
