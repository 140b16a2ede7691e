"""One encoder-shape attention forward (for ncu captures): python profiles/ubench/attn_fwd_one.py [causal] [long]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import ops  # noqa: E402

causal = "causal" in sys.argv
n, mean, H, d = (8, 4000, 12, 64) if "long" in sys.argv else (64, 995, 12, 64)
g = torch.Generator().manual_seed(1)
lens = (mean * (0.6 + 0.8 * torch.rand(n, generator=g))).long().clamp_(min=1)
cu = torch.zeros(n + 1, dtype=torch.int32)
cu[1:] = lens.cumsum(0)
cu = cu.cuda()
T, mx = int(lens.sum()), int(lens.max())
qkv = torch.randn(T, 3 * H * d, device="cuda").bfloat16()
q, k, v = qkv[:, :H * d], qkv[:, H * d:2 * H * d], qkv[:, 2 * H * d:]
for _ in range(4):
    out = ops.attn_varlen(q, k, v, cu, cu, mx, H, d, causal=causal)
torch.cuda.synchronize()
