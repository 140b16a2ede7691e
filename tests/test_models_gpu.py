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


def test_mae_batched_random_masking_invariants():
    """The no-`noises` path draws all images' masks at once (batched argsort): per image the kept ids, the restore
    permutation and the loss mask must be mutually consistent exactly as mask_sequence (models.py:106-125) defines them."""
    from acai_omr_b200 import synth
    torch.manual_seed(3)
    enc = M.MAEEncoder(0.75, 16, 12, 40, num_layers=1, num_heads=2, hidden_dim=64, mlp_dim=128).to(DEV).eval()
    imgs = [synth.synth_system(h, w, i).to(DEV) for i, (h, w) in enumerate([(64, 160), (96, 320), (32, 640), (80, 176)])]
    with torch.no_grad():
        h, pk, info = enc.embed_masked_packed(imgs)
    assert h.shape[0] == sum(info["kept_seq_lens"]) == pk.total
    seen = set()
    for L, lk, keep_mask, restore in zip(info["unmasked_seq_lens"], info["kept_seq_lens"], info["seq_masks"], info["ids_restore"]):
        assert lk == int(L * 0.25) and restore.shape == (L,) and keep_mask.shape == (L,)
        r = restore.cpu()
        assert sorted(r.tolist()) == list(range(L))                       # a permutation of the patch positions
        assert torch.equal(keep_mask.cpu() == 0, r < lk)                  # patch j is kept iff it lands in the first lk slots
        assert int(keep_mask.sum()) == L - lk
        seen.add(tuple(r[:8].tolist()))
    assert len(seen) > 1                                                  # different images, different draws
    with torch.no_grad():                                                 # a second draw differs from the first
        _, _, info2 = enc.embed_masked_packed(imgs)
    assert not torch.equal(info2["ids_restore"][1], info["ids_restore"][1])
