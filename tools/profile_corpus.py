"""configs[4] on one GPU (1M x 768 fp32 corpus, 32 queries, top-10) inside a cudaProfiler range."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200

g = torch.Generator(device="cuda").manual_seed(3)
corpus = torch.randn(1_000_000, 768, device="cuda", generator=g)
shard = lms_b200.CorpusShard(corpus)
qs = torch.randn(32, 768, device="cuda", generator=g)
for _ in range(3):
    shard.topk(qs, 10)
torch.cuda.synchronize()
torch.cuda.profiler.start()
shard.topk(qs, 10)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled one corpus search")
