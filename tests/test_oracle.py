"""The oracle (oracle/vitomr_oracle.py) against the reference's own outputs and known answers.

Golden entries were produced by the unmodified reference (tests/golden/make_golden.py).
CPU only; this is the pin that lets the GPU tests trust the oracle."""
import os

import pytest
import torch

from oracle import vitomr_oracle as O

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

TOL = dict(atol=2e-5, rtol=1e-4)


def sub(sd, prefix):
    return {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}


def test_encoder_batchify_and_forward(golden):
    g = golden["encoder"]
    cfg = dict(g["cfg"], stacks=["encoder_blocks"])
    emb, mask = O.encoder_batchify(g["imgs"], g["sd"], 4, 6, 10)
    assert torch.equal(mask, g["mask"])
    assert torch.allclose(emb, g["batchify"], **TOL)
    y, _ = O.encoder_forward(g["imgs"], g["sd"], cfg, fast_path_pad_fill=True)
    assert torch.allclose(y, g["out_eval"], **TOL)  # includes pad rows == final-norm bias
    y_t, _ = O.encoder_forward(g["imgs"], g["sd"], cfg, fast_path_pad_fill=False)
    valid = ~g["mask"]
    assert torch.allclose(y_t[valid], g["out_train"][valid], **TOL)
    y1, _ = O.encoder_forward([g["imgs"][1]], g["sd"], cfg)
    assert torch.allclose(y1, g["out_single"], **TOL)


def test_finetune_encoder_with_interpolated_pe(golden):
    g = golden["finetune_encoder"]
    assert torch.allclose(O.interpolate_pe(g["sd"]["pos_embedding"], 6, 3), g["pe_interp_6x3"], atol=1e-6)
    cfg = dict(patch_size=4, pe_max_height=4, pe_max_width=6, num_heads=2, interpolate=True,
               stacks=["frozen_blocks", "fine_tune_blocks"])
    emb, mask = O.encoder_batchify(g["imgs"], g["sd"], 4, 4, 6, interpolate=True)
    assert torch.equal(mask, g["mask"])
    assert torch.allclose(emb, g["batchify"], **TOL)
    y, _ = O.encoder_forward(g["imgs"], g["sd"], cfg)
    assert torch.allclose(y, g["out_eval"], **TOL)


def test_decoder_paths(golden):
    g = golden["decoder"]
    sd, H = g["sd"], g["cfg"]["num_heads"]
    tf = O.decoder_forward(g["tokens"], g["memory"], sd, H, g["lmx_mask"], g["mem_mask"])
    assert torch.allclose(tf, g["tf_logits"], **TOL)
    gen = O.decoder_forward(g["tokens"], g["memory"], sd, H, None, g["mem_mask"])
    assert torch.allclose(gen, g["gen_logits"], **TOL)
    st = O.prepare_caches(g["memory"], sd, H)
    assert torch.allclose(st.k_cross[0], g["k_cross0"], **TOL)
    assert torch.allclose(st.v_cross[0], g["v_cross0"], **TOL)
    steps = [O.cached_generate(g["tokens"][:, t:t + 1], t, st, sd, H, g["mem_mask"]).squeeze(1)
             for t in range(g["tokens"].shape[1])]
    assert torch.allclose(torch.stack(steps, 1), g["cached_logits"], **TOL)
    assert torch.allclose(st.k_self[2], g["k_self2"], **TOL)
    try:
        O.cached_generate(g["tokens"][:, :1], 24, st, sd, H)
    except RuntimeError:
        pass
    else:
        raise AssertionError("time_step >= max_lmx_seq_len must raise (models.py:519-520)")


def test_vitomr_greedy_and_teacher_forcing(golden):
    g = golden["vitomr"]
    sd = g["sd"]
    enc_cfg = dict(patch_size=4, pe_max_height=6, pe_max_width=10, num_heads=4, interpolate=True,
                   stacks=["fine_tune_blocks"])
    lat, lat_mask = O.encoder_forward(g["imgs"], sub(sd, "encoder."), enc_cfg)
    lat = O.transition_head(lat, sd)
    assert torch.equal(lat_mask, g["latent_mask"])
    assert torch.allclose(lat, g["latent"], **TOL)
    seqs, logp, m, _ = O.cached_greedy_generate(lat, sub(sd, "decoder."), 4, lat_mask, max_len=20)
    assert torch.equal(seqs, g["seqs"])          # token-exact, incl. the time_step=t quirk
    assert torch.equal(m, g["seq_mask"])
    assert torch.allclose(logp, g["log_probs"], **TOL)
    # a fixture that exercises both early-stopping rows and never-stopping rows
    assert bool((g["seqs"] == 2).any()) and not bool((g["seqs"] == 2).any(dim=1).all())
    cfg = dict(encoder=enc_cfg, decoder_heads=4, pad_idx=1)
    pred, tgt = O.teacher_forced_forward(list(zip(g["imgs"], g["lmx"])), sd, cfg)
    assert torch.equal(tgt, g["tf_target"])
    assert torch.allclose(pred, g["tf_pred"], **TOL)
    assert torch.allclose(O.ce_loss(pred, tgt, 1), g["tf_loss"], **TOL)


def test_mae_forward_and_loss(golden):
    g = golden["mae"]
    cfg = dict(mask_ratio=0.75, patch_size=4, pe_max_height=6, pe_max_width=10, encoder_heads=4, decoder_heads=2)
    pred, lmask, target = O.mae_forward(list(zip(g["imgs"], g["targets"])), g["sd"], cfg, g["noises"])
    assert torch.equal(lmask, g["loss_mask"])
    assert torch.equal(target, g["target"])
    valid = lmask  # masked & non-pad rows are what the loss reads
    assert torch.allclose(pred[valid], g["pred"][valid], **TOL)
    assert torch.allclose(O.mae_loss(pred, lmask, target), g["loss"], **TOL)


def test_mae_loss_known_answer():
    """Reference tests/test_mae.py:169-180."""
    target = torch.cat([
        torch.tensor([[1, 1, 1], [2, 2, 2]], dtype=torch.float).unsqueeze(-1).repeat(1, 1, 6),
        torch.tensor([[2, 2, 2], [3, 3, 3]], dtype=torch.float).unsqueeze(-1).repeat(1, 1, 6)], dim=-1)
    pred = torch.tensor([[2, 2, 2], [3, 3, 4]], dtype=torch.float).unsqueeze(-1).repeat(1, 1, 12)
    loss_mask = torch.tensor([[1, 0, 0], [1, 0, 1]], dtype=torch.float)
    assert O.mae_loss(pred, loss_mask, target) == 10.583329200744629


def test_batchify_known_answers():
    """Reference tests/test_vitomr.py:27-49 (identity projection, all-ones PE, incl. interpolated PE)."""
    sd = {"pos_embedding": torch.ones(50, 50, 4)}
    x = [torch.ones(1, 4, 4), torch.ones(1, 4, 8)]
    emb, mask = O.encoder_batchify(x, sd, 2, 60, 200, interpolate=True, identity_projection=True)
    first = torch.cat([torch.ones(1, 4, 4) + 1, torch.zeros(1, 4, 4)], dim=1)
    assert torch.equal(emb, torch.cat([first, torch.ones(1, 8, 4) + 1]))
    assert torch.equal(mask, torch.cat([(torch.arange(8) >= 4)[None], (torch.arange(8) >= 8)[None]]))
    sd = {"pos_embedding": torch.ones(60, 200, 4)}
    emb, _ = O.encoder_batchify([torch.ones(1, 10, 600)], sd, 2, 60, 200, interpolate=True, identity_projection=True)
    assert torch.equal(emb, torch.ones(1, 1500, 4) + 1)


def test_split_lmx_known_answer():
    """Reference tests/test_vitomr.py:151-172."""
    seqs = [torch.tensor([0, 2, 3, 226]), torch.tensor([0, 2, 2, 3, 4, 226])]
    inp, tgt, mask = O.batchify_and_split_lmx_seqs(seqs, 1)
    assert torch.equal(inp, torch.tensor([[0, 2, 3, 226, 1], [0, 2, 2, 3, 4]]))
    assert torch.equal(tgt, torch.tensor([[2, 3, 226, 1, 1], [2, 2, 3, 4, 226]]))
    assert torch.equal(mask, torch.tensor([[False, False, False, False, True], [False] * 5]))


def _leaf_sd(sd):
    return {k: v.clone().requires_grad_(True) if v.is_floating_point() else v for k, v in sd.items()}


def _check_grads(sd, ref, tol):
    for k, g_ref in ref.items():
        got = sd[k].grad
        assert got is not None, k
        assert torch.allclose(got, g_ref, **tol), (k, float((got - g_ref).abs().max()))


def test_teacher_forced_gradients_match_reference(golden, golden_grads):
    """d(OMRCELoss)/d(every parameter) of the oracle's autograd vs the reference's (train mode, dropout 0)."""
    g, gg = golden["vitomr"], golden_grads["vitomr"]
    sd = _leaf_sd(g["sd"])
    enc_cfg = dict(patch_size=4, pe_max_height=6, pe_max_width=10, num_heads=4, interpolate=True, stacks=["fine_tune_blocks"])
    pred, tgt = O.teacher_forced_forward(list(zip(g["imgs"], g["lmx"])), sd, dict(encoder=enc_cfg, decoder_heads=4, pad_idx=1))
    loss = O.ce_loss(pred, tgt, 1)
    assert torch.allclose(loss.detach(), gg["loss"], **TOL)
    loss.backward()
    _check_grads(sd, gg["grads"], dict(atol=2e-5, rtol=1e-3))


def test_mae_gradients_match_reference(golden, golden_grads):
    g, gg = golden["mae"], golden_grads["mae"]
    sd = _leaf_sd(g["sd"])
    cfg = dict(mask_ratio=0.75, patch_size=4, pe_max_height=6, pe_max_width=10, encoder_heads=4, decoder_heads=2)
    pred, lmask, target = O.mae_forward(list(zip(g["imgs"], g["targets"])), sd, cfg, g["noises"])
    loss = O.mae_loss(pred, lmask, target)
    assert torch.allclose(loss.detach(), gg["loss"], **TOL)
    loss.backward()
    _check_grads(sd, gg["grads"], dict(atol=2e-5, rtol=1e-3))


def test_scheduled_sampling_forward_and_gradients_match_reference(golden, golden_grads):
    """ScheduledSamplingViTOMR.forward_train + OMRCELoss + backward, soft and hard (straight-through) sampling, with the
    reference's own random draws replayed (tests/golden/make_golden_grads.py)."""
    g = golden["vitomr"]
    enc_cfg = dict(patch_size=4, pe_max_height=6, pe_max_width=10, num_heads=4, interpolate=True, stacks=["fine_tune_blocks"])
    cfg = dict(encoder=enc_cfg, decoder_heads=4, pad_idx=1)
    for mode in ("soft", "hard"):
        gg = golden_grads["scheduled_sampling"][mode]
        sd = _leaf_sd(g["sd"])
        pred, tgt = O.scheduled_sampling_forward(list(zip(g["imgs"], g["lmx"])), sd, cfg, gg["tf_prob"], gg["tau"], mode == "hard",
                                                 gg["rand"], gg["gumbel"])
        assert torch.equal(tgt, gg["target"])
        valid = tgt != 1
        assert torch.allclose(pred[valid], gg["pred"][valid], **TOL)
        loss = O.ce_loss(pred, tgt, 1)
        assert torch.allclose(loss.detach(), gg["loss"], **TOL)
        loss.backward()
        _check_grads(sd, gg["grads"], dict(atol=2e-5, rtol=1e-3))


# ------------------------------------------------------------------------------------------ GRPO
def test_grpo_objective_restatements_match_reference(golden_grpo):
    """oracle.grpo_objective / entropy_bonus and the product-side acai_omr_b200.grpo functions against the values the
    reference's own calc_grpo_objective / calc_entropy_bonus produced, including the gradient w.r.t. the logits."""
    from acai_omr_b200 import grpo
    g = golden_grpo
    c = g["cfg"]
    V = g["theta_logits"].shape[-1]
    obj = O.grpo_objective(g["theta_logits"], g["rollouts"], g["attn_mask"], g["log_probs"], g["advantages"], c["epsilon"], c["num_groups"])
    bonus = O.entropy_bonus(g["theta_logits"], g["attn_mask"], V)
    assert torch.allclose(obj, g["objective"], atol=1e-6, rtol=1e-5) and torch.allclose(bonus, g["bonus"], atol=1e-6, rtol=1e-5)
    z = g["theta_logits"].clone().requires_grad_(True)
    obj2 = grpo.calc_grpo_objective(z, g["rollouts"], g["attn_mask"], g["log_probs"], g["advantages"], c["epsilon"], c["num_groups"])
    bonus2 = grpo.calc_entropy_bonus(z, g["attn_mask"], V)
    (obj2 + c["entropy_beta"] * bonus2).backward()
    assert torch.allclose(obj2, g["objective"], atol=1e-6, rtol=1e-5) and torch.allclose(bonus2, g["bonus"], atol=1e-6, rtol=1e-5)
    assert torch.allclose(z.grad, g["dlogits"], atol=1e-7, rtol=1e-4)
    r = torch.tensor([[1.0, 2.0, 4.0], [0.5, 0.5, 2.0]])
    adv = grpo.group_advantages(r)
    assert adv.shape == (6,) and torch.allclose(adv.view(2, 3).mean(dim=1), torch.zeros(2), atol=1e-6)


def test_oracle_rollout_policy_matches_reference(golden_grpo):
    """The oracle's rollout (cached decoder step + written-out top-k / temperature / inverse-CDF draw) reproduces the
    unmodified reference's rollouts, per-token log-probs and masks for the recorded uniforms."""
    g = golden_grpo
    c = g["cfg"]
    sd = {k[len("decoder."):]: v for k, v in g["policy_sd"].items() if k.startswith("decoder.")}
    rollouts, logp, mask = O.cached_forward_rollout_policy(g["xlatent"], sd, 4, g["uniforms"], g["xmask"], max_actions=c["max_actions"],
                                                           top_k=c["top_k"], temperature=c["temperature"])
    assert torch.equal(rollouts, g["rollouts"]) and torch.equal(mask, g["rollout_mask"])
    assert torch.allclose(logp, g["log_probs"], atol=2e-5, rtol=1e-4)


# ---- input pipeline: DynamicResize (utils/utils.py:334-370) ----------------------------------------------------------
def test_oracle_dynamic_resize_matches_reference_golden():
    """The explicit weight-matrix restatement of torch's antialiased bicubic against outputs of the reference's own
    DynamicResize class (tests/golden/make_golden_resize.py): down / up-scaling, portrait, uint8 and fp32, centre crops."""
    entries = torch.load(os.path.join(GOLDEN_DIR, "golden_resize.pt"))
    assert len(entries) >= 9
    for e in entries:
        out = O.dynamic_resize(e["img"], *e["args"])
        assert out.shape == e["out"].shape and out.dtype == torch.float32
        assert float((out - e["out"]).abs().max()) <= 2e-6
        assert float(out.min()) >= 0.0 and float(out.max()) <= 1.0


def test_oracle_dynamic_resize_target_sizes():
    """DynamicResize(16, N, 60, 200) shapes of SURVEY section 8(d): N = 1024 and N = 512 mixes."""
    for (h, w), want in {(512, 512): (512, 512), (300, 650): (352, 704), (100, 310): (288, 864), (90, 400): (256, 1024),
                         (80, 640): (176, 1408)}.items():
        assert O.dynamic_resize_target(h, w, 16, 1024) == want
    assert O.dynamic_resize_target(700, 300, 16, 512) == (512, 256)        # portrait: aspect = height // width
    assert O.dynamic_resize_target(100, 420, 16, 512) == (176, 704)


# ---- bf16 inference mode (vitomr_inference.py:79-86,94) pinned to the unmodified reference's own bf16 run ------------
@pytest.mark.parametrize("name", ["flat", "peaked"])
def test_oracle_bf16_mode_matches_reference_recorded_run(name):
    """tests/golden/golden_bf16.pt = the reference's transition head + cached greedy loop under
    torch.autocast("cpu", bfloat16) with bf16 KV caches (make_golden_bf16.py).  The oracle's `autocast_bf16()` mode,
    driven by the reference's tokens, must give the reference's per-step logits within the bf16 gate (2e-2; measured
    8e-3 / 1.1e-2: one bf16 ulp of the largest logit - torch's CPU bf16 kernels round partial sums differently), agree on
    every argmax except inside the tie band, and every free-running divergence must sit at a tie."""
    import os
    from oracle import parity as P
    g = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_bf16.pt"),
                   map_location="cpu", weights_only=False)[name]
    sd, H, max_len = g["sd"], g["cfg"]["num_heads"], g["cfg"]["max_len"]
    dec = {k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}
    with torch.no_grad(), O.autocast_bf16():
        head = O.transition_head(g["latent"], sd)
        forced = O.cached_forced_logits(head, dec, H, g["seqs"], g["mask"])
        seqs, logp, m, _ = O.cached_greedy_generate(head, dec, H, g["mask"], max_len=max_len)
    assert head.dtype == torch.bfloat16 and forced.dtype == torch.bfloat16
    n = forced.shape[1]
    valid = g["seq_mask"][:, :n]
    assert P.errors(head, g["head"], ~g["mask"])["max_rel"] < 4e-3
    e = P.errors(forced, g["logits"][:, :n], valid)
    assert e["max_rel"] < P.BF16_TOL and e["l2_rel"] < P.BF16_TOL, e
    a = P.argmax_agreement(forced, g["logits"][:, :n], P.BF16_TOL, valid)
    assert a["not_tie_disagreements"] == 0 and a["agree"] > 0.9, a
    d = P.divergence_report(seqs, g["seqs"], g["logits"], P.BF16_TOL)
    assert d["exact_or_tie_flip_rate"] == 1.0, d
    # the fp32 oracle sits inside the same gate: the 2e-2 tolerance is about bf16 rounding, not about the algorithm
    with torch.no_grad():
        forced32 = O.cached_forced_logits(O.transition_head(g["latent"], sd), dec, H, g["seqs"], g["mask"])
    assert P.errors(forced32, g["logits"][:, :n], valid)["max_rel"] < P.BF16_TOL
