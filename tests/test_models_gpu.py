"""GPU: host-side serving helpers around the drop-in modules (acai_omr_b200/pipeline.py)."""
import pytest
import torch

from acai_omr_b200.models import models as M

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_encoder_pipeline_matches_synchronous_calls():
    """EncoderPipeline.run (overlapped H2D / compute / D2H) returns exactly what Encoder.forward returns, batch by batch."""
    from acai_omr_b200 import synth
    from acai_omr_b200.pipeline import EncoderPipeline
    torch.manual_seed(0)
    enc = M.FineTuneOMREncoder(16, 60, 200, 2, num_layers=3, hidden_dim=64, num_heads=2, mlp_dim=128).to(DEV).eval()
    batches = []
    for b in range(5):
        shapes = [(64 + 16 * ((b + i) % 3), 160 + 32 * ((2 * b + i) % 4)) for i in range(3 + b % 2)]
        batches.append([synth.synth_system(h, w, 100 * b + i).pin_memory() for i, (h, w) in enumerate(shapes)])
    want = []
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        for batch in batches:
            out, mask = enc([t.to(DEV) for t in batch])
            want.append((out.cpu(), mask.cpu()))
    pipe = EncoderPipeline(enc, DEV, depth=2)
    n = 0
    for (out_h, mask_h), (out_w, mask_w) in zip(pipe.run(iter(batches)), want):
        assert out_h.is_pinned() and torch.equal(out_h, out_w) and torch.equal(mask_h, mask_w)
        n += 1
    assert n == len(batches)
    assert list(pipe.run([])) == []
