"""One BERT-base relevance batch (configs[1]: 32 x 128-token queries, 32 x 512-token texts, bf16) inside a
cudaProfiler range for ncu --profile-from-start off."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200
from oracle import reference_path as rp

dtype = sys.argv[1] if len(sys.argv) > 1 else "bf16"
bert = lms_b200.B200BertModel.from_hf(rp.build_bert(seed=0), dtype=dtype)
q, a = rp.bert_ids(32, 128, seed=2).cuda(), rp.bert_ids(32, 512, seed=5).cuda()
for _ in range(2):
    bert.embed(q); bert.embed(a)
torch.cuda.synchronize()
torch.cuda.profiler.start()
bert.embed(a)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled one 32x512 BERT forward")
