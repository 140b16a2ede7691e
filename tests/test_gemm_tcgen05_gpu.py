"""GPU: the tcgen05/TMA GEMM (bf16 operands, K % 64 == 0) against fp32 torch on the same bf16-rounded operands.
Covers every N-tile width the dispatcher picks, M/N tails, all epilogues and persistent multi-tile CTAs."""
import pytest
import torch

from acai_omr_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


def check(M, N, K, act=ops.ACT_NONE, res_dtype=None, out_dtype=torch.bfloat16, bias=True):
    x = rnd(M, K, seed=M + N).bfloat16()
    w = rnd(N, K, seed=K, scale=K ** -0.5).bfloat16()
    b = rnd(N, seed=3) if bias else None
    r = rnd(M, N, seed=4).to(res_dtype) if res_dtype is not None else None
    y = ops.linear(x, w, b, act=act, residual=r, out_dtype=out_dtype)
    ref = x.float() @ w.float().t()
    if bias:
        ref = ref + b
    if act == ops.ACT_GELU:
        ref = torch.nn.functional.gelu(ref)
    if r is not None:
        ref = ref + r.float()
    assert y.dtype == out_dtype and y.shape == (M, N)
    err = (y.float() - ref).abs().max().item()
    tol = 2e-2 if out_dtype == torch.bfloat16 else 2e-3
    assert torch.allclose(y.float(), ref, rtol=tol, atol=tol), f"M={M} N={N} K={K}: max abs err {err}"


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (128, 256, 128), (1, 64, 64), (200, 768, 768), (333, 2304, 768),
                                   (1000, 3072, 768), (700, 768, 3072), (256, 1024, 1024), (256, 4096, 1024),
                                   (64, 227, 1024), (20000, 768, 256), (5, 8, 64)])
def test_tcgen05_gemm_shapes(M, N, K):
    check(M, N, K)


def test_tcgen05_gemm_large_persistent():
    check(40000, 2304, 768)  # 313 x 9 tiles of 128x256 over 148 persistent CTAs
    check(40000, 768, 3072, res_dtype=torch.bfloat16)


@pytest.mark.parametrize("act", [ops.ACT_NONE, ops.ACT_GELU])
@pytest.mark.parametrize("res_dtype", [None, torch.float32, torch.bfloat16])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_tcgen05_gemm_epilogues(act, res_dtype, out_dtype):
    check(300, 384, 192, act=act, res_dtype=res_dtype, out_dtype=out_dtype)
    check(130, 227, 128, act=act, res_dtype=res_dtype, out_dtype=out_dtype, bias=False)  # scalar-store tail path


def test_tcgen05_gemm_strided_operands():
    qkv = rnd(500, 3 * 256, seed=1).bfloat16()
    w = rnd(3 * 128, 256, seed=2, scale=0.06).bfloat16()
    y = ops.linear(qkv[:, 256:512], w[128:256], None, out_dtype=torch.float32)
    ref = qkv[:, 256:512].float() @ w[128:256].float().t()
    assert torch.allclose(y, ref, rtol=2e-3, atol=2e-3)
