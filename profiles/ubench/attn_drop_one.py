"""One encoder-shape attention forward + backward WITH attention dropout (for ncu captures of the DROP kernel variants):
    python profiles/ubench/attn_drop_one.py [p]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import ops  # noqa: E402

p = float(sys.argv[1]) if len(sys.argv) > 1 else 0.05
n, mean, H, d = 64, 995, 12, 64
g = torch.Generator().manual_seed(1)
lens = (mean * (0.6 + 0.8 * torch.rand(n, generator=g))).long().clamp_(min=1)
cu = torch.zeros(n + 1, dtype=torch.int32)
cu[1:] = lens.cumsum(0)
cu = cu.cuda()
T, mx = int(lens.sum()), int(lens.max())
qkv = torch.randn(T, 3 * H * d, device="cuda").bfloat16()
q, k, v = qkv[:, :H * d], qkv[:, H * d:2 * H * d], qkv[:, 2 * H * d:]
dqkv = torch.empty_like(qkv)
for _ in range(3):
    out, lse = ops.attn_varlen(q, k, v, cu, cu, mx, H, d, causal=False, return_lse=True, dropout_p=p, dropout_seed=1234)
    dout = torch.randn_like(out)
    ops.attn_varlen_bwd(q, k, v, out, dout, lse, cu, cu, mx, mx, H, d, False, dqkv[:, :H * d], dqkv[:, H * d:2 * H * d],
                        dqkv[:, 2 * H * d:], dropout_p=p, dropout_seed=1234)
torch.cuda.synchronize()
