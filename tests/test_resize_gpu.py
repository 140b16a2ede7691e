"""GPU: acai_dynamic_resize (DynamicResize, utils/utils.py:334-370, + ToImage / ToDtype(scale=True) for uint8 input) through
the drop-in `transforms.DynamicResize` against the reference-generated golden and the CPU oracle."""
import os

import pytest
import torch

from acai_omr_b200 import transforms
from acai_omr_b200._lib import AcaiError
from oracle import vitomr_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_resize.pt")
TOL = 3e-6       # fp32 summation order only (the reference's own output is the target)


def test_dynamic_resize_matches_reference_golden():
    for e in torch.load(GOLDEN):
        mod = transforms.DynamicResize(*e["args"])
        out = mod(e["img"].to(DEV))
        assert out.shape == e["out"].shape and out.dtype == torch.float32 and out.is_cuda
        err = float((out.cpu() - e["out"]).abs().max())
        assert err <= TOL, (tuple(e["img"].shape), e["args"], err)


def test_dynamic_resize_ragged_batch_equals_single_calls_and_oracle():
    g = torch.Generator().manual_seed(3)
    shapes = [(64, 300), (200, 90), (33, 33), (120, 1000), (50, 51)]
    imgs = [torch.randint(0, 256, (1, h, w), generator=g, dtype=torch.uint8) if i % 2 == 0 else torch.rand(1, h, w, generator=g)
            for i, (h, w) in enumerate(shapes)]
    mod = transforms.DynamicResize(16, 128, 60, 200, True)
    outs = mod.forward_batch([t.to(DEV) for t in imgs])
    assert len(outs) == len(imgs)
    for t, o in zip(imgs, outs):
        ref = O.dynamic_resize(t, 16, 128, 60, 200, True)
        assert o.shape == ref.shape
        assert float((o.cpu() - ref).abs().max()) <= TOL
        assert torch.equal(o, mod(t.to(DEV)))                      # batched = one at a time, bit for bit
        assert o.shape[-2] % 16 == 0 and o.shape[-1] % 16 == 0     # what Encoder.batchify needs


def test_dynamic_resize_uint8_equals_scaled_float_and_2d_input():
    g = torch.Generator().manual_seed(4)
    u8 = torch.randint(0, 256, (1, 77, 410), generator=g, dtype=torch.uint8)
    mod = transforms.DynamicResize(16, 256, 60, 200, False)
    a = mod(u8.to(DEV))
    b = mod((u8.float() / 255.0).to(DEV))
    assert float((a - b).abs().max()) <= 1e-6
    assert torch.equal(mod(u8[0].to(DEV)), a)                       # (H, W) is accepted like (1, H, W)
    strided = torch.zeros(77, 500, dtype=torch.uint8)
    strided[:, :410] = u8[0]
    assert torch.equal(mod(strided.to(DEV)[:, :410]), a)            # row pitch != width


def test_dynamic_resize_feeds_the_encoder():
    from acai_omr_b200.models import models as M
    torch.manual_seed(0)
    enc = M.OMREncoder(16, 12, 40, num_layers=1, hidden_dim=64, num_heads=2, mlp_dim=128).to(DEV).eval()
    g = torch.Generator().manual_seed(5)
    scans = [torch.randint(0, 256, (1, 90, 700), generator=g, dtype=torch.uint8).to(DEV),
             torch.randint(0, 256, (1, 200, 210), generator=g, dtype=torch.uint8).to(DEV)]
    imgs = transforms.DynamicResize(16, 64, 12, 40, True).forward_batch(scans)
    with torch.no_grad():
        lat, mask = enc(imgs)
    assert lat.shape[0] == 2 and lat.shape[1] == max((i.shape[-2] // 16) * (i.shape[-1] // 16) for i in imgs)
    assert bool(torch.isfinite(lat).all())


def test_dynamic_resize_errors():
    mod = transforms.DynamicResize(16, 64, 60, 200, False)
    with pytest.raises(AcaiError):
        mod(torch.rand(1, 32, 32))                                   # CPU tensor: no fallback
    with pytest.raises(AcaiError):
        mod(torch.rand(3, 32, 32, device=DEV))                       # NUM_CHANNELS = 1
    with pytest.raises(AcaiError):
        mod(torch.rand(1, 32, 32, device=DEV).double())
    assert transforms.resize_batch([], 16, 64, 60, 200, False) == []
