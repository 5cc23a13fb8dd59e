"""The decode-attention kernel alone at the configs[2] mid-point shape (B=64, H=12, S=640, bf16), rotating over 12
layers of cache, for `ncu --set full -k regex:attn_decode`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import lms_b200

lib, check = lms_b200._lib.load(), lms_b200._lib.check
B, H, S, L, Smax = 64, 12, 640, 12, 768
g = torch.Generator(device="cuda").manual_seed(0)
k = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
v = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
q = torch.randn(B, H * 64, device="cuda", generator=g).bfloat16()
out = torch.empty_like(q)
for rep in range(3):
    for l in range(L):
        check(lib.tutor_attn_decode(1, q.data_ptr(), k[l].data_ptr(), v[l].data_ptr(), out.data_ptr(), B, H, Smax, S, None))
torch.cuda.synchronize()
print("ran", 3 * L, "attention launches")
