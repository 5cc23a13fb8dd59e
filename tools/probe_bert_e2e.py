"""Where does the host-path BERT relevance batch (configs[1]) spend its time?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lms_b200
from oracle import reference_path as rp
bert = lms_b200.B200BertModel.from_hf(rp.build_bert(seed=0), dtype="bf16")
q, a = rp.bert_ids(32, 128, seed=2), rp.bert_ids(32, 512, seed=5)
qd, ad = q.cuda(), a.cuda()
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
eq, ea = bert.embed(q), bert.embed(a)
eqd, ead = bert.embed(qd), bert.embed(ad)
print(f"embed(q host) {t(lambda: bert.embed(q)):.2f} ms   embed(q dev) {t(lambda: bert.embed(qd)):.2f} ms")
print(f"embed(a host) {t(lambda: bert.embed(a)):.2f} ms   embed(a dev) {t(lambda: bert.embed(ad)):.2f} ms")
print(f"cosine host {t(lambda: lms_b200.cosine_similarity_batch(eq, ea)):.2f} ms   cosine dev {t(lambda: lms_b200.cosine_similarity_batch(eqd, ead)):.2f} ms")
print(f"gate host {t(lambda: lms_b200.cosine_similarity_batch(bert.embed(q), bert.embed(a))):.2f} ms   gate dev {t(lambda: lms_b200.cosine_similarity_batch(bert.embed(qd), bert.embed(ad))):.2f} ms")
