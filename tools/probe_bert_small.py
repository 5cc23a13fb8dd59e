"""BERT-base embed latency for the product's real request shapes (one query of <= 128 tokens)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
hf = rp.build_bert(seed=0)
for dtype in ("bf16",):
    bert = lms_b200.B200BertModel.from_hf(hf, dtype=dtype)
    for B, T in ((1, 16), (1, 48), (1, 100), (1, 128), (2, 128)):
        ids = rp.bert_ids(B, T, seed=2)
        for _ in range(3): bert.embed(ids)
        t0 = time.perf_counter(); n = 20
        for _ in range(n): bert.embed(ids)
        print(f"{dtype} B={B} T={T}: {(time.perf_counter() - t0) / n * 1e3:.3f} ms per embed (host ids in, host embedding out)", flush=True)
    del bert
