"""Isolated timing of the fused GEMM + residual + LayerNorm cluster kernel against the two-launch path, encoder shapes.
    python profiles/ubench/linear_ln_bench.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import _lib, ops  # noqa: E402

dev = "cuda"
lib = _lib.load()
print("resident clusters:", {cl: lib.acai_linear_layernorm_resident_clusters(cl) for cl in (2, 3, 4)})
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


def back_to_back(fns, reps=10):
    """mean us per call of a chain queued without events in between (what the launch gaps cost)"""
    for f in fns:
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for f in fns:
            f()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


g = torch.Generator().manual_seed(0)
SHAPES = [(63664, 768, 768), (63664, 768, 3072), (41976, 1024, 1024), (41976, 1024, 4096)]
if os.environ.get("LNB_SHAPE"):
    SHAPES = [tuple(int(v) for v in os.environ["LNB_SHAPE"].split(","))]
for (M, N, K) in SHAPES:
    x = torch.randn(M, K, generator=g).bfloat16().to(dev)
    w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16().to(dev)
    b, ga, be = (torch.randn(N, generator=g).to(dev) for _ in range(3))
    for rdt in (torch.float32, torch.bfloat16):
        r = torch.randn(M, N, generator=g).to(rdt).to(dev)
        want32 = rdt == torch.float32
        fused = lambda: ops.linear_layernorm(x, w, b, r, ga, be, 1e-5, want_f32=want32)

        def two():
            y = ops.linear(x, w, b, residual=r, out_dtype=rdt)
            return ops.layernorm(y, ga, be, 1e-5, out_dtype=rdt, out2_dtype=torch.bfloat16 if want32 else None)
        gemm_only = lambda: ops.linear(x, w, b, residual=r, out_dtype=rdt)
        tf, tt, tg = timeit(fused), timeit(two), timeit(gemm_only)
        byts = M * K * 2 + N * K * 2 + M * N * (r.element_size() + 2 + (4 if want32 else 0))
        print(f"M={M} N={N} K={K} res={str(rdt)[6:]}: fused {tf:7.1f} us ({2.0 * M * N * K / tf / 1e6:6.0f} TFLOP/s, {byts / tf / 1e3:5.0f} GB/s)  "
              f"two-launch {tt:7.1f} us (gemm alone {tg:7.1f})  chain of 4: fused {back_to_back([fused] * 4) / 4:7.1f}  two {back_to_back([two] * 4) / 4:7.1f} us/call")
