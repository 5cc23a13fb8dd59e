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
s, i = shard.search(qd, k)
torch.cuda.synchronize()
dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    s, i = shard.search(qd, k)
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / 5], device="cuda")
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    g = torch.Generator().manual_seed(3)
    full = torch.cat([torch.randn(min(100_000, N - st), 768, generator=g) for st in range(0, N, 100_000)])
    rs, ri = rp.corpus_topk(queries, full, k)
    ok = torch.equal(i.cpu().long(), ri) and (s.cpu() - rs).abs().max().item() < 1e-4
    print(f"sharded search over {world} GPUs: {'OK' if ok else 'MISMATCH'}; {ms.item() * 1e3:.1f} us per 32-query search "
          f"(max over ranks), {N * 768 * 4 / world / (ms.item() * 1e-3) / 1e9:.0f} GB/s per GPU")
    assert ok
dist.destroy_process_group()
