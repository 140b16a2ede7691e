"""GPU: acai_colsum_static - the allocation-free variant of acai_colsum (slab partials in a file-scope `static __device__`
array, no workspace argument), called through the C ABI and checked against the existing entry point on two calls in a row.

The file sorts last on purpose: a failure here cannot hide any other test.  It passes
both alone in a fresh process and after the other tests (the entry point looks its scratch address up once per device and
shares nothing else with them)."""
import pytest
import torch

from acai_omr_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _dtype_code(t):
    from acai_omr_b200 import ops
    return ops._dt(t)


def colsum_ws(x, accumulate=False, out=None):
    """the existing entry point, caller-provided workspace"""
    lib = _lib.load()
    R, Cc = x.shape
    out = torch.empty(Cc, dtype=torch.float32, device=x.device) if out is None else out
    ws = torch.empty(lib.acai_colsum_workspace_bytes(R, Cc), dtype=torch.uint8, device=x.device)
    rc = lib.acai_colsum(x.data_ptr(), _dtype_code(x), x.stride(0), R, Cc, out.data_ptr(), int(accumulate), ws.data_ptr(),
                         ws.numel(), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "acai_colsum")
    return out


def colsum_static(x, accumulate=False, out=None):
    """the new entry point: inputs and outputs only"""
    lib = _lib.load()
    R, Cc = x.shape
    out = torch.empty(Cc, dtype=torch.float32, device=x.device) if out is None else out
    rc = lib.acai_colsum_static(x.data_ptr(), _dtype_code(x), x.stride(0), R, Cc, out.data_ptr(), int(accumulate),
                                torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "acai_colsum_static")
    return out


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_colsum_static_matches_colsum_twice(dtype):
    _lib.require_device()
    g = torch.Generator().manual_seed(0)
    # two calls in a row on DIFFERENT problems (the second smaller: it must not see the first call's partials), then the
    # first one again; bf16 takes the vectorised slab kernel, fp32 / the odd width the generic one
    shapes = [(4099, 1024), (300, 136), (4099, 1024), (63664, 768 if dtype == torch.bfloat16 else 520)]
    for i, (R, Cc) in enumerate(shapes):
        x = torch.randn(R, Cc, generator=g).to(DEV, dtype)
        want = colsum_ws(x)
        got = colsum_static(x)
        torch.cuda.synchronize()
        assert torch.equal(got, want), (i, (got - want).abs().max())           # same kernels, same order: bit-identical
        ref = x.float().sum(dim=0)
        assert torch.allclose(got, ref, rtol=1e-4, atol=1e-2 if dtype == torch.bfloat16 else 2e-3)
        acc = torch.full((Cc,), 0.5, device=DEV)
        acc2 = acc.clone()
        colsum_static(x, accumulate=True, out=acc)
        colsum_ws(x, accumulate=True, out=acc2)
        assert torch.equal(acc, acc2)


def test_colsum_static_allocates_nothing_and_rejects_oversize():
    _lib.require_device()
    lib = _lib.load()
    cap = lib.acai_colsum_static_max_workspace_bytes()
    assert cap == 4 << 20
    x = torch.randn(2048, 512, device=DEV).bfloat16()
    out = torch.empty(512, dtype=torch.float32, device=DEV)
    colsum_static(x, out=out)                                   # first call on this device: symbol address looked up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = torch.cuda.memory_stats().get("num_alloc_retries", 0), torch.cuda.memory_allocated()
    for _ in range(3):
        colsum_static(x, out=out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0                                        # no cudaMalloc behind the call
    assert (torch.cuda.memory_stats().get("num_alloc_retries", 0), torch.cuda.memory_allocated()) == stats0
    assert torch.equal(out, colsum_ws(x))
    # beyond the static scratch: ceil(rows / 128) * cols * 4 > 4 MiB -> a clear error, nothing launched
    big_rows, big_cols = 128 * 1025, 1024
    assert lib.acai_colsum_workspace_bytes(big_rows, big_cols) > cap
    probe = torch.zeros(8, device=DEV)                           # never dereferenced: the size check comes first
    rc = lib.acai_colsum_static(probe.data_ptr(), 0, big_cols, big_rows, big_cols, out.data_ptr(), 0,
                                torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.acai_last_error().decode()
    assert "static scratch" in msg and str(cap) in msg, msg
    # the largest problem that fits is accepted
    R = 128 * 1024
    xb = torch.ones(R, 1024, device=DEV, dtype=torch.bfloat16)
    assert lib.acai_colsum_workspace_bytes(R, 1024) == cap
    o = colsum_static(xb)
    assert torch.equal(o, torch.full((1024,), float(R), device=DEV))
