"""us per acai_linear_fwd launch for decode-step shapes (bf16), M sweep.  python profiles/ubench/linear_small_m.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import ops

torch.manual_seed(0)
for N, K in [(1024, 1024), (3072, 1024), (4096, 1024), (1024, 4096)]:
    ws = [torch.randn(N, K, device="cuda").bfloat16() for _ in range(12)]      # 12 layers' worth: no L2 reuse of weights
    b = torch.randn(N, device="cuda")
    row = []
    for M in (1, 2, 4, 8, 16, 32, 33, 64, 128, 256):
        x = torch.randn(M, K, device="cuda").bfloat16()
        for w in ws:
            ops.linear(x, w, b)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()                      # replayed graph: device time per launch, no Python in the way
        with torch.cuda.graph(g):
            for _ in range(5):
                for w in ws:
                    ops.linear(x, w, b)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        row.append(f"M={M}: {e0.elapsed_time(e1) / 60 * 1e3:.1f}")
    print(f"N={N} K={K}  us/launch  " + "  ".join(row), flush=True)
