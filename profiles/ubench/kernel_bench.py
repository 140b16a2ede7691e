"""Isolated timings (CUDA events, L2 flushed between launches) of the non-GEMM kernels of the encoder / training step at
the bench's shapes: attention forward / backward, LayerNorm forward, column sums.  Used for A/B runs through the
library's environment switches (e.g. ACAI_ATTN_POLY=0|1|2).  Usage: python profiles/ubench/kernel_bench.py [what ...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import ops  # noqa: E402

DEV = "cuda"
FLUSH = None


def timed(fn, iters=10, warmup=3):
    global FLUSH
    if FLUSH is None:
        FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(warmup):
        fn()
    tot = 0.0
    for _ in range(iters):
        FLUSH.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3          # us


def ragged(n, mean, seed):
    g = torch.Generator().manual_seed(seed)
    lens = (mean * (0.6 + 0.8 * torch.rand(n, generator=g))).long().clamp_(min=1)
    cu = torch.zeros(n + 1, dtype=torch.int32)
    cu[1:] = lens.cumsum(0)
    return lens, cu.to(DEV)


def bench_attn(name, n, mean, H, d, causal, bwd, drop=0.0):
    lens, cu = ragged(n, mean, 1)
    T, mx = int(lens.sum()), int(lens.max())
    qkv = torch.randn(T, 3 * H * d, device=DEV).bfloat16()
    q, k, v = qkv[:, :H * d], qkv[:, H * d:2 * H * d], qkv[:, 2 * H * d:]
    flops = float(4 * H * d * (lens.double() ** 2).sum()) * (0.5 if causal else 1.0)
    dk_ = dict(dropout_p=drop, dropout_seed=1234) if drop else {}
    if drop:
        name = f"{name} p={drop}"
    us = timed(lambda: ops.attn_varlen(q, k, v, cu, cu, mx, H, d, causal=causal, **dk_))
    print(f"{name:28s} fwd  {us:9.1f} us  {flops / us / 1e6:7.1f} TFLOP/s  (T={T})")
    if bwd:
        out, lse = ops.attn_varlen(q, k, v, cu, cu, mx, H, d, causal=causal, return_lse=True, **dk_)
        dout = torch.randn_like(out)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv[:, :H * d], dqkv[:, H * d:2 * H * d], dqkv[:, 2 * H * d:]
        us = timed(lambda: ops.attn_varlen_bwd(q, k, v, out, dout, lse, cu, cu, mx, mx, H, d, causal, dq, dk, dv, **dk_))
        print(f"{name:28s} bwd  {us:9.1f} us  {2.5 * flops / us / 1e6:7.1f} TFLOP/s")


def bench_cross(name, n, mean_q, mean_k, H, d, drop=0.0):
    """decoder cross-attention: queries = target tokens, keys = encoder memory (fwd + bwd)"""
    lq, cu_q = ragged(n, mean_q, 3)
    lk, cu_k = ragged(n, mean_k, 1)
    Tq, Tk, mq, mk = int(lq.sum()), int(lk.sum()), int(lq.max()), int(lk.max())
    q = torch.randn(Tq, H * d, device=DEV).bfloat16()
    kv = torch.randn(Tk, 2 * H * d, device=DEV).bfloat16()
    k, v = kv[:, :H * d], kv[:, H * d:]
    flops = float(4 * H * d * (lq.double() * lk.double()).sum())
    dk_ = dict(dropout_p=drop, dropout_seed=1234) if drop else {}
    us = timed(lambda: ops.attn_varlen(q, k, v, cu_q, cu_k, mq, H, d, causal=False, **dk_))
    print(f"{name:28s} fwd  {us:9.1f} us  {flops / us / 1e6:7.1f} TFLOP/s  (Tq={Tq}, Tk={Tk})")
    out, lse = ops.attn_varlen(q, k, v, cu_q, cu_k, mq, H, d, causal=False, return_lse=True, **dk_)
    dout = torch.randn_like(out)
    dq, dkv = torch.empty_like(q), torch.empty_like(kv)
    us = timed(lambda: ops.attn_varlen_bwd(q, k, v, out, dout, lse, cu_q, cu_k, mq, mk, H, d, False, dq, dkv[:, :H * d], dkv[:, H * d:], **dk_))
    print(f"{name:28s} bwd  {us:9.1f} us  {2.5 * flops / us / 1e6:7.1f} TFLOP/s")


def bench_ln(rows, cols):
    x = torch.randn(rows, cols, device=DEV).bfloat16()
    g, b = torch.ones(cols, device=DEV), torch.zeros(cols, device=DEV)
    us = timed(lambda: ops.layernorm(x, g, b, 1e-5, out_dtype=torch.bfloat16))
    print(f"layernorm bf16 {rows}x{cols}: {us:8.1f} us  {4.0 * rows * cols / us / 1e3:7.1f} GB/s")


def bench_colsum(rows, cols):
    x = torch.randn(rows, cols, device=DEV).bfloat16()
    out = torch.zeros(cols, device=DEV)
    us = timed(lambda: ops.colsum(x, out=out, accumulate=True))
    print(f"colsum bf16 {rows}x{cols}: {us:8.1f} us  {2.0 * rows * cols / us / 1e3:7.1f} GB/s")


def bench_resize(n=64):
    """64 uint8 scans at ~1.7x the DynamicResize(16, 1024, 60, 200) target of their aspect ratio -> fp32 encoder inputs."""
    from acai_omr_b200 import synth, transforms
    g = torch.Generator().manual_seed(2)
    scans = []
    for h, w in synth.system_shapes(n):
        H, W = int(h * 1.7) + 3, int(w * 1.7) + 5
        scans.append(torch.randint(0, 256, (1, H, W), generator=g, dtype=torch.uint8).to(DEV))
    mod = transforms.DynamicResize(16, 1024, 60, 200, False)
    outs = mod.forward_batch(scans)
    nbytes = sum(s_.numel() for s_ in scans) + 4 * sum(o.numel() for o in outs)
    us = timed(lambda: mod.forward_batch(scans))
    print(f"dynamic resize, {n} uint8 scans ({sum(s_.numel() for s_ in scans) / 1e6:.1f} MB) -> fp32 ({sum(o.numel() for o in outs) * 4 / 1e6:.1f} MB): "
          f"{us:8.1f} us  {n / us * 1e6:9.0f} systems/s  {nbytes / us / 1e3:7.1f} GB/s (src + dst)")


if __name__ == "__main__":
    what = sys.argv[1:] or ["attn", "ln", "colsum"]
    print("ACAI_ATTN_POLY =", os.environ.get("ACAI_ATTN_POLY", "(default)"), " ACAI_ATTN_CTAS =", os.environ.get("ACAI_ATTN_CTAS", "(default)"))
    if "attn" in what or "attn_bwd" in what:
        bwd = "attn_bwd" in what
        bench_attn("encoder self (12 x 64)", 64, 995, 12, 64, False, bwd)
        bench_attn("decoder self (16 x 64, causal)", 64, 655, 16, 64, True, bwd)
        bench_attn("MAE decoder self (16 x 32)", 64, 495, 16, 32, False, bwd)
        if bwd:
            bench_cross("decoder cross (16 x 64)", 64, 655, 995, 16, 64)
        if "drop" in what:
            bench_attn("encoder self (12 x 64)", 64, 995, 12, 64, False, bwd, 0.05)
            bench_attn("decoder self (16 x 64, causal)", 64, 655, 16, 64, True, bwd, 0.1)
    if "ln" in what:
        bench_ln(63664, 768)
        bench_ln(41976, 1024)
    if "resize" in what:
        bench_resize()
    if "colsum" in what:
        bench_colsum(63664, 768)
        bench_colsum(63664, 3072)
        bench_colsum(41976, 1024)
