"""Relative error (vs fp64) of the tcgen05 weight-gradient kernel as a function of the token-range split rule:
    ACAI_WGRAD_SPLIT=0|1 python profiles/ubench/wgrad_precision.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import ops  # noqa: E402

torch.manual_seed(0)
for T, n_out, k_in in [(20000, 1024, 4096), (63664, 768, 3072), (63664, 3072, 768), (41976, 1024, 4096), (63664, 768, 768), (41976, 1024, 1024)]:
    dz = torch.randn(T, n_out, device="cuda").bfloat16()
    x = torch.randn(T, k_in, device="cuda").bfloat16()
    dw = torch.zeros(n_out, k_in, device="cuda")
    ops.wgrad_accumulate(dz, x, dw)
    ref = dz[:, :256].double().t() @ x.double()
    err = float((dw[:256].double() - ref).norm() / ref.norm())
    mx = float((dw[:256].double() - ref).abs().max() / ref.abs().max())
    # the same products summed by torch in fp32 (cuBLAS bf16 GEMM with fp32 accumulation), for scale
    tb = (dz[:, :256].t().contiguous() @ x).float()
    terr = float((tb.double() - ref).norm() / ref.norm())
    print(f"split rule {os.environ.get('ACAI_WGRAD_SPLIT', '1')}  T={T} {n_out}x{k_in}: rel l2 {err:.2e}  max-norm {mx:.2e}   (torch bf16 matmul -> bf16 out: {terr:.2e})")
