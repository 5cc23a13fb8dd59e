"""configs[1] (32 x 128-token queries, 32 x 512-token texts, BERT-base): bf16 engine time per batch (CUDA events) and its
cosine / embedding deviation from the fp32 engine.  TUTOR_BERT_F32_STREAM=1 selects the fp32 activation stream."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200
from oracle import reference_path as rp

hf = rp.build_bert(seed=0)
q, a = rp.bert_ids(32, 128, seed=2).cuda(), rp.bert_ids(32, 512, seed=5).cuda()
f32 = lms_b200.B200BertModel.from_hf(hf, dtype="f32")
ref_q, ref_a = f32.embed(q).float().cpu(), f32.embed(a).float().cpu()
ref_cos = torch.nn.functional.cosine_similarity(ref_q, ref_a)
del f32
bert = lms_b200.B200BertModel.from_hf(hf, dtype="bf16")
for _ in range(3):
    eq, ea = bert.embed(q), bert.embed(a)
torch.cuda.synchronize()
cos = torch.nn.functional.cosine_similarity(eq.float().cpu(), ea.float().cpu())
print(f"stream={'f32' if os.environ.get('TUTOR_BERT_F32_STREAM') == '1' else 'bf16'}  "
      f"cosine max abs err {float((cos - ref_cos).abs().max()):.3e}  "
      f"embedding max abs err {float((ea.float().cpu() - ref_a).abs().max()):.3e}")
n = 20
for name, fn in (("32x128 + 32x512", lambda: (bert.embed(q), bert.embed(a))), ("32x512", lambda: bert.embed(a))):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"  {name}: {e0.elapsed_time(e1) / n:.3f} ms")
