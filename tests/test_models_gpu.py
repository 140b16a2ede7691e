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


def test_mask_sequence_reference_test():
    """reference tests/test_mae.py:36-55 (MAEEncoder.mask_sequence): labelled patches survive shuffle -> mask -> unshuffle."""
    from torch import nn
    enc = M.MAEEncoder(0.50, 2, 2, 4, num_heads=1, hidden_dim=1)
    L = 4
    x = torch.arange(L).unsqueeze(0).repeat(12, 1).unsqueeze(0)           # (1, C*P*P = 12, L), column j holds j
    pe_num_grid = torch.arange(4, dtype=torch.float).reshape(2, 2)
    enc.pos_embedding = nn.Parameter(torch.cat((pe_num_grid, torch.zeros(2, 2) - 1), dim=1).unsqueeze(-1))
    t_masked, pe_slice, unmasked_len, len_keep, seq_mask, ids_restore = enc.mask_sequence(x, 2, 2)
    assert t_masked.shape == torch.Size([1, 12, 2]) and unmasked_len == L and len_keep == 2
    assert pe_slice.shape == (2, 1) and torch.equal(pe_slice.squeeze(1), t_masked[0, 0].float())    # PE follows its patch
    assert seq_mask.sum() == L - len_keep
    undo = torch.cat((t_masked, torch.zeros(1, 12, 2, dtype=t_masked.dtype) - 1), dim=-1).index_select(-1, ids_restore)
    assert undo.shape == x.shape
    kept = seq_mask == 0
    assert torch.equal(undo[0, 0][kept], x[0, 0][kept]) and bool((undo[0, 0][~kept] == -1).all())
    # injected noise: argsort order decides what is kept
    t2, _, _, _, m2, _ = enc.mask_sequence(x, 2, 2, noise=torch.tensor([0.9, 0.1, 0.5, 0.2]))
    assert t2[0, 0].tolist() == [1, 3] and m2.tolist() == [1, 0, 1, 0]


def test_in_place_edit_of_the_latent_is_honoured():
    """The padded latent carries a note of the packed rows it was made from; an in-place edit must invalidate the note so
    that the transition head / decoder see what the caller wrote, as the reference would."""
    torch.manual_seed(0)
    enc = M.FineTuneOMREncoder(16, 12, 40, 1, num_layers=1, hidden_dim=64, num_heads=2, mlp_dim=128)
    dec = M.OMRDecoder(16, M.default_vocab_path(), num_layers=1, hidden_dim=64, num_heads=2, mlp_dim=128)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=64).to(DEV).eval()
    imgs = [torch.rand(1, 32, 64).to(DEV), torch.rand(1, 48, 48).to(DEV)]
    tok = torch.randint(3, 227, (2, 5)).to(DEV)
    with torch.no_grad():
        lat, mask = vit.encoder(imgs)
        base = vit.decoder.generate(tok, vit.transition_head(lat), latent_attention_mask=mask)
        lat2, mask2 = vit.encoder(imgs)
        lat2.mul_(0.5)                                         # in-place edit between encoder and head
        edited = vit.decoder.generate(tok, vit.transition_head(lat2), latent_attention_mask=mask2)
        want = vit.decoder.generate(tok, vit.transition_head(lat.clone() * 0.5), latent_attention_mask=mask)
        assert torch.allclose(edited, want, atol=1e-5) and not torch.allclose(edited, base, atol=1e-3)
        # the mask too: hiding more of the memory must change the result
        lat3, mask3 = vit.encoder(imgs)
        head3 = vit.transition_head(lat3)
        mask3[0, 4:] = True
        hidden = vit.decoder.generate(tok, head3, latent_attention_mask=mask3)
        assert not torch.allclose(hidden[0], base[0], atol=1e-4) and torch.allclose(hidden[1], base[1], atol=1e-5)


def test_teacher_forced_generate_and_checkpoint_grads():
    """TeacherForcedViTOMR.generate (models.py:749-757) and OMRDecoder.forward(checkpoint_grads=True) (models.py:470-478):
    checkpointing changes memory, not numbers - same logits, same gradients, with dropout on."""
    from acai_omr_b200 import ops
    torch.manual_seed(0)
    enc = M.FineTuneOMREncoder(16, 12, 40, 1, num_layers=1, hidden_dim=64, num_heads=2, mlp_dim=128)
    dec = M.OMRDecoder(16, M.default_vocab_path(), num_layers=3, hidden_dim=64, num_heads=2, mlp_dim=128, transformer_dropout=0.1)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=64).to(DEV)
    vit.eval()
    with torch.no_grad():
        lat, _ = vit.encoder(torch.rand(1, 32, 64).to(DEV))
        head = vit.transition_head(lat)
        seqs = torch.randint(3, 227, (4, 6)).to(DEV)
        nxt = vit.generate(head, seqs)
        ref = torch.log_softmax(vit.decoder.generate(seqs, head.expand(4, -1, -1))[:, -1].float(), dim=-1)
    assert nxt.shape == (4, 227) and torch.allclose(nxt, ref, atol=1e-6) and torch.allclose(nxt.exp().sum(-1), torch.ones(4, device=DEV), atol=1e-5)
    vit.train()
    tok = torch.randint(3, 227, (2, 7)).to(DEV)
    mem = torch.rand(2, 9, 64, device=DEV)
    lm = torch.zeros(2, 7, dtype=torch.bool, device=DEV)
    mm = torch.zeros(2, 9, dtype=torch.bool, device=DEV)
    grads = []
    for ck in (False, True):
        ops.DropoutState.manual_seed(123)
        vit.zero_grad()
        out = vit.decoder(tok, mem, lm, mm, checkpoint_grads=ck)
        out.float().square().mean().backward()
        grads.append((out.detach().clone(), [p.grad.clone() for p in vit.decoder.parameters() if p.grad is not None]))
    assert torch.equal(grads[0][0], grads[1][0])
    assert len(grads[0][1]) == len(grads[1][1]) > 10
    for a, b in zip(grads[0][1], grads[1][1]):
        assert torch.allclose(a, b, atol=1e-6, rtol=1e-4)
