"""Corpus scan + top-k time on one GPU over shard sizes (the per-GPU work of the sharded search at 1/2/4/8 GPUs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
g = torch.Generator(device="cuda").manual_seed(3)
qs = torch.randn(32, 768, device="cuda", generator=g)
for N in ([int(a) for a in sys.argv[1:]] or [1_000_000, 500_000, 250_000, 125_000]):
    corpus = torch.randn(N, 768, device="cuda", generator=g)
    shard = lms_b200.CorpusShard(corpus)
    shard.topk(qs, 10); torch.cuda.synchronize()
    side = torch.cuda.Stream(); graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            shard.topk(qs, 10)
        graph.replay(); side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(20): graph.replay()
        e1.record(side); side.synchronize()
    us = e0.elapsed_time(e1) * 50
    print(f"N={N}: {us:.1f} us per search, {N * 768 * 4 / us / 1e3:.0f} GB/s ({N * 768 * 4 / us / 1e3 / 6549.1:.3f} of HBM peak)", flush=True)
    del corpus, shard
