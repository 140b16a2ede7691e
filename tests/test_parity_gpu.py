"""GPU: the drop-in modules (CUDA path, through the C ABI) against the reference's golden outputs
and against the CPU oracle on seeded inputs.  fp32 mode: logits within 1e-3 relative (BASELINE.json
north_star) - in practice ~1e-5 - and greedy tokens exact."""
import pytest
import torch

from acai_omr_b200.models import models as M
from oracle import vitomr_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
VOCAB = M.default_vocab_path()
TOL = dict(atol=5e-5, rtol=1e-4)


def cuda_list(ts):
    return [t.to(DEV) for t in ts]


def rel_err(a, b):
    """max-norm relative error; `P.errors` (oracle/parity.py) also gives the l2 and per-element figures, used by the bf16 and
    full-size tests."""
    return ((a.float().cpu() - b.float().cpu()).abs().max() / b.float().abs().max().clamp_min(1e-6)).item()


def sub(sd, prefix):
    return {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}


def test_encoder_matches_reference_golden(golden):
    g = golden["encoder"]
    kw = {k: g["cfg"][k] for k in ("num_layers", "hidden_dim", "num_heads", "mlp_dim")}
    enc = M.Encoder(4, 6, 10, **kw)
    enc.load_state_dict(g["sd"])
    enc = enc.to(DEV).eval()
    with torch.no_grad():
        emb, mask = enc.batchify(cuda_list(g["imgs"]))
        y, mask2 = enc(cuda_list(g["imgs"]))
        y1 = enc.generate(g["imgs"][1].to(DEV))
    assert torch.equal(mask.cpu(), g["mask"]) and torch.equal(mask2.cpu(), g["mask"])
    assert torch.allclose(emb.cpu(), g["batchify"], **TOL)       # incl. padding rows = projection bias
    assert torch.allclose(y.cpu(), g["out_eval"], **TOL)         # incl. padding rows = final-norm bias
    assert torch.allclose(y1.cpu(), g["out_single"], **TOL)
    enc.train()  # Encoder has dropout 0.0: deterministic slow path in the reference
    with torch.no_grad():
        yt, _ = enc(cuda_list(g["imgs"]))
    valid = ~g["mask"]
    assert torch.allclose(yt.cpu()[valid], g["out_train"][valid], **TOL)


def test_finetune_encoder_interpolated_pe_golden(golden):
    g = golden["finetune_encoder"]
    c = g["cfg"]
    enc = M.FineTuneOMREncoder(4, 4, 6, c["fine_tune_depth"], num_layers=c["num_layers"], hidden_dim=c["hidden_dim"],
                               num_heads=c["num_heads"], mlp_dim=c["mlp_dim"], transformer_dropout=0.0)
    enc.load_state_dict(g["sd"])
    enc = enc.to(DEV).eval()
    with torch.no_grad():
        pe = enc.interpolate_pe(6, 3)
        emb, _ = enc.batchify(cuda_list(g["imgs"]))
        y, mask = enc(cuda_list(g["imgs"]))
    assert torch.allclose(pe.cpu(), g["pe_interp_6x3"], atol=1e-6)
    assert torch.allclose(emb.cpu(), g["batchify"], **TOL)
    assert torch.equal(mask.cpu(), g["mask"])
    assert torch.allclose(y.cpu(), g["out_eval"], **TOL)


def test_decoder_paths_golden(golden):
    g = golden["decoder"]
    c = g["cfg"]
    kw = dict(num_layers=c["num_layers"], hidden_dim=c["hidden_dim"], num_heads=c["num_heads"], mlp_dim=c["mlp_dim"])
    dec_u = M.OMRDecoder(24, VOCAB, **kw)
    dec_c = M.OMRDecoder(24, VOCAB, **kw, use_caching=True, max_batch_size=4, cache_dtype=torch.float)
    dec_u.load_state_dict(g["sd"])
    dec_c.load_state_dict(g["sd"])
    dec_u, dec_c = dec_u.to(DEV).eval(), dec_c.to(DEV).eval()
    mem, mm, tok, lm = g["memory"].to(DEV), g["mem_mask"].to(DEV), g["tokens"].to(DEV), g["lmx_mask"].to(DEV)
    with torch.no_grad():
        tf = dec_u(tok, mem, lm, mm)
        gen = dec_u.generate(tok, mem, latent_attention_mask=mm)
        dec_c.prepare_caches(mem)
        steps = torch.stack([dec_c.cached_generate(tok[:, t:t + 1], t, latent_attention_mask=mm).squeeze(1)
                             for t in range(tok.shape[1])], dim=1)
        k0, v0 = dec_c.decoder_blocks.cross_attn_caches[0].get_cached_keys_and_vals()
        ks2 = dec_c.decoder_blocks.self_attn_caches[2].k_cache[:3, :, :tok.shape[1]]
        gen_c = dec_c.generate(tok, mem, latent_attention_mask=mm)  # uncached path through the cached module
    valid = ~g["lmx_mask"]
    assert torch.allclose(tf.cpu()[valid], g["tf_logits"][valid], **TOL)
    assert torch.allclose(gen.cpu(), g["gen_logits"], **TOL)
    assert torch.allclose(gen_c.cpu(), g["gen_logits"], **TOL)
    assert torch.allclose(steps.cpu(), g["cached_logits"], **TOL)
    assert torch.allclose(k0.cpu(), g["k_cross0"], **TOL) and torch.allclose(v0.cpu(), g["v_cross0"], **TOL)
    assert torch.allclose(ks2.cpu(), g["k_self2"], **TOL)
    with pytest.raises(RuntimeError):
        dec_c.cached_generate(tok[:, :1], 24)


def build_vitomr(g):
    c = g["cfg"]
    enc = M.FineTuneOMREncoder(4, 6, 10, 2, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64)
    dec = M.OMRDecoder(20, VOCAB, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96, use_caching=True,
                       max_batch_size=4, cache_dtype=torch.float)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=c["transition_head_dim"])
    vit.load_state_dict(g["sd"])
    return vit.to(DEV).eval()


def test_vitomr_greedy_token_exact_golden(golden):
    g = golden["vitomr"]
    vit = build_vitomr(g)
    with torch.no_grad():  # as inference() does (vitomr_inference.py:73-86), fp32 throughout
        lat, lat_mask = vit.encoder(cuda_list(g["imgs"]))
        lat = vit.transition_head(lat)
        seqs, logp, smask = vit.cached_greedy_generate(lat, lat_mask, max_len=20)
    valid = ~g["latent_mask"]
    assert torch.allclose(lat.cpu()[valid], g["latent"][valid], **TOL)
    assert torch.equal(seqs.cpu(), g["seqs"])       # token-exact, mixed early-stop / never-stop rows
    assert torch.equal(smask.cpu(), g["seq_mask"])
    assert torch.allclose(logp.cpu(), g["log_probs"], **TOL)
    events = list(vit.streamed_cached_greedy_generate(lat[1:2], lat_mask[1:2], max_len=20, flush_interval=3))
    assert events[-1]["type"] == "inference_finish" and all(e["type"] == "step" for e in events[:-1])
    n = int(g["seq_mask"][1].sum())
    assert torch.equal(events[-1]["payload"]["sequence"].cpu()[0, :n], g["seqs"][1, :n])
    with pytest.raises(ValueError):
        next(vit.streamed_cached_greedy_generate(lat, lat_mask))


def test_teacher_forced_forward_golden(golden):
    g = golden["vitomr"]
    vit = build_vitomr(g)
    batch = list(zip(cuda_list(g["imgs"]), cuda_list(g["lmx"])))
    with torch.no_grad():
        pred, tgt = vit(batch)
    assert torch.equal(tgt.cpu(), g["tf_target"])
    inp, _, lmx_mask = O.batchify_and_split_lmx_seqs(g["lmx"], 1)
    valid = ~lmx_mask
    assert torch.allclose(pred.cpu()[valid], g["tf_pred"][valid], **TOL)
    assert torch.allclose(O.ce_loss(pred.cpu(), tgt.cpu(), 1), g["tf_loss"], **TOL)


def randomised_sd(module, seed):
    gen = torch.Generator().manual_seed(seed)
    sd = {}
    for k, v in module.state_dict().items():
        if "norm" in k and k.endswith("weight"):
            sd[k] = 1.0 + 0.02 * torch.randn(v.shape, generator=gen)
        elif k.endswith("bias"):
            sd[k] = 0.01 * torch.randn(v.shape, generator=gen)
        else:
            sd[k] = 0.05 * torch.randn(v.shape, generator=gen)
    return sd


def test_mid_size_pipeline_vs_oracle():
    """Seeded mid-size model (d_head 64, 16x16 patches, ragged 'system' shapes): CUDA vs CPU oracle."""
    enc = M.FineTuneOMREncoder(16, 12, 40, 2, num_layers=3, hidden_dim=128, num_heads=2, mlp_dim=256)
    dec = M.OMRDecoder(48, VOCAB, num_layers=2, hidden_dim=192, num_heads=3, mlp_dim=384, use_caching=True,
                       max_batch_size=4, cache_dtype=torch.float)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=160)
    sd = randomised_sd(vit, 21)
    vit.load_state_dict(sd)
    vit = vit.to(DEV).eval()
    g = torch.Generator().manual_seed(22)
    imgs = [torch.rand(1, h, w, generator=g) for h, w in [(64, 160), (32, 320), (96, 96), (16, 640)]]
    enc_cfg = dict(patch_size=16, pe_max_height=12, pe_max_width=40, num_heads=2, interpolate=True,
                   stacks=["frozen_blocks", "fine_tune_blocks"])
    lat_o, mask_o = O.encoder_forward(imgs, sub(sd, "encoder."), enc_cfg)
    lat_o = O.transition_head(lat_o, sd)
    seqs_o, logp_o, m_o, logits_o = O.cached_greedy_generate(lat_o, sub(sd, "decoder."), 3, mask_o, max_len=24)
    with torch.no_grad():
        lat, mask = vit.encoder(cuda_list(imgs))
        lat = vit.transition_head(lat)
        logits0 = vit.decoder.generate(seqs_o[:, :8].to(DEV), lat, latent_attention_mask=mask)
        seqs, logp, m = vit.cached_greedy_generate(lat, mask, max_len=24)
    assert torch.equal(mask.cpu(), mask_o)
    valid = ~mask_o
    assert rel_err(lat.cpu()[valid], lat_o[valid]) < 1e-3
    # padding rows too: the reference applies the head to them (all equal head(final-norm bias) in eval mode)
    assert mask_o.any() and torch.allclose(lat.cpu()[mask_o], lat_o[mask_o], atol=1e-5, rtol=1e-4)
    ref_logits = O.decoder_forward(seqs_o[:, :8], lat_o, sub(sd, "decoder."), 3, None, mask_o)
    assert rel_err(logits0, ref_logits) < 1e-3      # north_star fp32 gate
    assert torch.equal(seqs.cpu(), seqs_o) and torch.equal(m.cpu(), m_o)
    assert torch.allclose(logp.cpu(), logp_o, atol=1e-4, rtol=1e-3)


def test_bf16_autocast_mode_within_tolerance():
    """bf16 mode as the reference runs it (vitomr_inference.py:80-84): fp32 encoder outside autocast, head and
    decoder under autocast.  north_star gate: logits within 2e-2 relative."""
    enc = M.OMREncoder(16, 12, 40, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256)
    dec = M.OMRDecoder(32, VOCAB, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, use_caching=True,
                       max_batch_size=4, cache_dtype=torch.bfloat16)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=128)
    sd = randomised_sd(vit, 31)
    vit.load_state_dict(sd)
    vit = vit.to(DEV).eval()
    g = torch.Generator().manual_seed(32)
    imgs = [torch.rand(1, h, w, generator=g) for h, w in [(64, 160), (32, 320), (96, 96)]]
    enc_cfg = dict(patch_size=16, pe_max_height=12, pe_max_width=40, num_heads=2, interpolate=True, stacks=["encoder_blocks"])
    lat_o, mask_o = O.encoder_forward(imgs, sub(sd, "encoder."), enc_cfg)
    lat_o = O.transition_head(lat_o, sd)
    tok = torch.randint(3, 227, (3, 10), generator=g)
    ref = O.decoder_forward(tok, lat_o, sub(sd, "decoder."), 2, None, mask_o)
    with torch.no_grad():
        lat, mask = vit.encoder(cuda_list(imgs))
        with torch.autocast("cuda", dtype=torch.bfloat16):
            lat = vit.transition_head(lat)
            logits = vit.decoder.generate(tok.to(DEV), lat, latent_attention_mask=mask)
            vit.decoder.prepare_caches(lat)
            steps = torch.stack([vit.decoder.cached_generate(tok[:, t:t + 1].to(DEV), t, latent_attention_mask=mask).squeeze(1)
                                 for t in range(10)], dim=1)
    assert logits.dtype == torch.bfloat16
    from oracle import parity as P
    for got in (logits, steps):
        e = P.errors(got, ref)
        assert e["max_rel"] < P.BF16_TOL and e["l2_rel"] < P.BF16_TOL, e
        a = P.argmax_agreement(got, ref, P.BF16_TOL)          # token agreement, aware of the reference's top-1 / top-2 margin
        assert a["not_tie_disagreements"] == 0, a


def test_mae_forward_and_loss_golden(golden):
    """MAE.forward + MAELoss (models.py:249-288) with the reference's own torch.rand draws injected."""
    g = golden["mae"]
    c = g["cfg"]
    mae = M.MAE(c["mask_ratio"], c["patch_size"], c["pe_max_height"], c["pe_max_width"], c["encoder_hidden_dim"],
                c["decoder_hidden_dim"], c["encoder_kwargs"], c["decoder_kwargs"])
    mae.load_state_dict(g["sd"])
    mae = mae.to(DEV).eval()
    batch = list(zip(cuda_list(g["imgs"]), cuda_list(g["targets"])))
    with torch.no_grad():
        pred, lmask, target = mae(batch, noises=g["noises"])
        loss = M.MAELoss()(pred, lmask, target)
    assert torch.equal(lmask.cpu(), g["loss_mask"])
    assert torch.equal(target.cpu(), g["target"])
    valid = g["loss_mask"]
    assert torch.allclose(pred.cpu()[valid], g["pred"][valid], **TOL)
    assert torch.allclose(loss.cpu(), g["loss"], **TOL)
    # padded-API pieces the reference's tests exercise (tests/test_mae.py:57-122)
    with torch.no_grad():
        emb, enc_mask, dec_mask, keep, full, seq_masks, restore, dims = mae.encoder.batchify(cuda_list(g["imgs"]), noises=g["noises"])
    assert keep == [int(L * 0.25) for L in full] and dec_mask.shape[1] == max(full)
    assert [tuple(m.shape) for m in seq_masks.unbind()] == [(L,) for L in full]


def test_mae_loss_known_answer_gpu():
    """Reference tests/test_mae.py:169-180: MAELoss == 10.583329200744629."""
    target = torch.cat([
        torch.tensor([[1, 1, 1], [2, 2, 2]], dtype=torch.float).unsqueeze(-1).repeat(1, 1, 6),
        torch.tensor([[2, 2, 2], [3, 3, 3]], dtype=torch.float).unsqueeze(-1).repeat(1, 1, 6)], dim=-1).to(DEV)
    pred = torch.tensor([[2, 2, 2], [3, 3, 4]], dtype=torch.float).unsqueeze(-1).repeat(1, 1, 12).to(DEV)
    loss_mask = torch.tensor([[1, 0, 0], [1, 0, 1]], dtype=torch.float).to(DEV)
    loss = M.MAELoss()(pred, loss_mask, target)
    assert abs(loss.item() - 10.583329200744629) < 2e-5


def test_mae_bf16_full_head_dims():
    """MAE decoder shape class of the real model: 16 heads of 32 (SIMT attention path), bf16 autocast vs the fp32 oracle."""
    torch.manual_seed(0)
    from acai_omr_b200 import synth
    mae = M.MAE(0.75, 16, 12, 40, encoder_hidden_dim=128, decoder_hidden_dim=64,
                encoder_kwargs=dict(num_layers=2, num_heads=2, mlp_dim=256), decoder_kwargs=dict(num_layers=2, num_heads=2, mlp_dim=128))
    synth.rerandomise_(mae, 3)
    sd = {k: v.detach().clone() for k, v in mae.state_dict().items()}
    g = torch.Generator().manual_seed(1)
    imgs = [torch.rand(1, 64, 160, generator=g), torch.rand(1, 32, 320, generator=g), torch.rand(1, 48, 48, generator=g)]
    noises = [torch.rand((t.shape[1] // 16) * (t.shape[2] // 16), generator=g) for t in imgs]
    cfg = dict(mask_ratio=0.75, patch_size=16, pe_max_height=12, pe_max_width=40, encoder_heads=2, decoder_heads=2)
    pred_o, lmask_o, target_o = O.mae_forward(list(zip(imgs, imgs)), sd, cfg, noises)
    mae = mae.to(DEV).eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        pred, lmask, target = mae(list(zip(cuda_list(imgs), cuda_list(imgs))), noises=noises)
        loss = M.MAELoss()(pred, lmask, target)
    assert torch.equal(lmask.cpu(), lmask_o) and torch.allclose(target.cpu(), target_o)
    assert rel_err(pred[lmask], pred_o[lmask_o]) < 2e-2
    assert abs(loss.item() - O.mae_loss(pred_o, lmask_o, target_o).item()) < 2e-2 * max(1.0, loss.item())
