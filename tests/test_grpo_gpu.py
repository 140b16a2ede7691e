"""GPU: the GRPO pieces of the hot path - device-side top-k sampling, GRPOViTOMR rollouts on both decode
paths, and the clipped-ratio objective's gradients - against the unmodified reference's golden (tests/golden/golden_grpo.pt)
and the CPU oracle."""
import math

import pytest
import torch

from acai_omr_b200 import grpo, ops
from acai_omr_b200.models import models as M
from oracle import vitomr_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
VOCAB = M.default_vocab_path()


def build_parts(c, max_batch):
    enc = M.FineTuneOMREncoder(4, 6, 10, 1, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64)
    dec = M.OMRDecoder(20, VOCAB, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96, use_caching=True, max_batch_size=max_batch,
                       cache_dtype=torch.float)
    head = M.TeacherForcedViTOMR(M.FineTuneOMREncoder(4, 6, 10, 1, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64), None,
                                 M.OMRDecoder(20, VOCAB, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96),
                                 transition_head_dim=40).transition_head
    return enc, head, dec


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,vocab,top_k,temperature", [(9, 227, 50, 1.2), (64, 227, 5, 0.7), (5, 33, 33, 1.0), (3, 1000, 1, 2.0),
                                                          (17, 64, 70, 1.0)])
def test_topk_sample_update_kernel_vs_oracle(dtype, rows, vocab, top_k, temperature):
    g = torch.Generator().manual_seed(rows * 1000 + vocab)
    logits = (3.0 * torch.randn(rows, vocab, generator=g)).to(dtype)
    logits[0, : min(8, vocab)] = logits[0, 0]                            # ties: the lower index wins a place in the top k
    steps = 6
    u = torch.rand(rows, steps, generator=g)
    u[1 % rows, 2] = 0.0
    u[2 % rows, 2] = 0.99999994
    seqs = torch.full((rows, steps), -1, dtype=torch.long, device=DEV)
    logp = torch.zeros(rows, steps, device=DEV)
    fin = torch.zeros(rows, dtype=torch.uint8, device=DEV)
    step = torch.zeros(1, dtype=torch.int32, device=DEV)
    for t in range(1, steps):
        step.fill_(t)
        ops.topk_sample_update(logits.to(DEV), step, seqs, logp, fin, 2, top_k, temperature, uniforms=u.to(DEV))
        tok, lp = O.topk_sample_step(logits.float(), top_k, temperature, u[:, t])
        assert torch.equal(seqs[:, t].cpu(), tok), t
        assert torch.allclose(logp[:, t].cpu(), lp, atol=2e-5, rtol=1e-5), t
    assert torch.equal(fin.bool().cpu(), (seqs[:, 1:].cpu() == 2).any(dim=1))


def test_topk_sample_update_hashed_stream_statistics():
    """Without explicit uniforms the draw comes from a hash of (seed, row, step): reproducible per seed, different across
    seeds / steps, and distributed as softmax(top-k logits / T) (chi-square over 64k rows sharing one logits row)."""
    rows, vocab, top_k, T = 65536, 227, 8, 1.3
    g = torch.Generator().manual_seed(5)
    row = 2.0 * torch.randn(vocab, generator=g)
    logits = row.repeat(rows, 1).to(DEV)
    seqs = torch.zeros((rows, 4), dtype=torch.long, device=DEV)
    logp = torch.zeros(rows, 4, device=DEV)
    fin = torch.zeros(rows, dtype=torch.uint8, device=DEV)
    step = torch.ones(1, dtype=torch.int32, device=DEV)
    seed = torch.full((1,), 1234, dtype=torch.int64, device=DEV)
    ops.topk_sample_update(logits, step, seqs, logp, fin, 2, top_k, T, seed=seed)
    first, first_lp = seqs[:, 1].clone(), logp[:, 1].clone()
    ops.topk_sample_update(logits, step, seqs, logp, fin, 2, top_k, T, seed=seed)
    assert torch.equal(seqs[:, 1], first)                                 # same (seed, row, step) -> same draw
    step.fill_(2)
    ops.topk_sample_update(logits, step, seqs, logp, fin, 2, top_k, T, seed=seed)
    assert not torch.equal(seqs[:, 2], first)
    seed.fill_(99)
    step.fill_(1)
    ops.topk_sample_update(logits, step, seqs, logp, fin, 2, top_k, T, seed=seed)
    assert not torch.equal(seqs[:, 1], first)
    kept = torch.topk(row, top_k).indices
    p = torch.softmax(row[kept] / T, dim=0)
    counts = torch.bincount(first.cpu(), minlength=vocab).float()
    assert int(counts.sum() - counts[kept].sum()) == 0                    # nothing outside the top k is ever drawn
    chi2 = float(((counts[kept] - rows * p) ** 2 / (rows * p)).sum())
    assert chi2 < 40.0, chi2                                             # 7 degrees of freedom: P(chi2 > 40) ~ 1e-6
    want_lp = torch.log_softmax(row[kept], dim=0)                        # recorded log-probs are temperature-free
    got = first_lp.cpu()
    for j, tok in enumerate(kept.tolist()):
        sel = first.cpu() == tok
        if bool(sel.any()):
            assert torch.allclose(got[sel], want_lp[j].expand(int(sel.sum())), atol=1e-5)


def test_grpo_vitomr_rollouts_match_reference_golden(golden_grpo):
    """GRPOViTOMR built from the teacher-forced state dict (frozen + fine-tuned encoder layers re-indexed), rollouts
    through the module-level cached path (head_dim 12: not the engine's shape) with the reference's recorded draws."""
    g = golden_grpo
    c = g["cfg"]
    enc, head, dec = build_parts(c, 12)
    policy = M.GRPOViTOMR(enc, head, dec, g["tf_sd"]).to(DEV).eval()
    assert isinstance(policy.encoder, M.OMREncoder) and not isinstance(policy.encoder, M.FineTuneOMREncoder)
    for k, v in policy.state_dict().items():
        assert torch.equal(v.cpu(), g["policy_sd"][k]), k
    assert all(not p.requires_grad for p in policy.encoder.parameters()) and all(not p.requires_grad for p in policy.transition_head.parameters())
    assert all(p.requires_grad for p in policy.decoder.parameters())
    with torch.no_grad():
        lat, lat_mask = policy.encoder([t.to(DEV) for t in g["imgs"]])
        lat = policy.transition_head(lat)
        valid = ~g["latent_mask"]                                          # padding rows are don't-care (masked as keys)
        assert torch.equal(lat_mask.cpu(), g["latent_mask"])
        assert torch.allclose(lat.cpu()[valid], g["latent"][valid], atol=5e-5, rtol=1e-4)
        xlat, xmask = policy.expand_img_latent_for_rollout(lat, lat_mask, c["group_size"])
        assert torch.equal(xmask.cpu(), g["xmask"])
        assert torch.allclose(xlat.cpu()[~g["xmask"]], g["xlatent"][~g["xmask"]], atol=5e-5, rtol=1e-4)
        rollouts, logp, rmask = policy.cached_forward_rollout_policy(xlat, xmask, max_actions=c["max_actions"], top_k=c["top_k"],
                                                                    temperature=c["temperature"], uniforms=g["uniforms"])
    assert torch.equal(rollouts.cpu(), g["rollouts"]) and torch.equal(rmask.cpu(), g["rollout_mask"])
    assert torch.allclose(logp.cpu(), g["log_probs"], atol=5e-5, rtol=1e-4)
    shifted, attn = policy.prepare_rollouts_for_policy_theta(rollouts, rmask)
    assert torch.equal(shifted.cpu(), g["shifted"]) and torch.equal(attn.cpu(), g["attn_mask"])


def test_grpo_objective_gradients_match_reference_golden(golden_grpo):
    """policy_theta's decoder forward on the right-shifted rollouts, clipped-ratio objective + entropy bonus, backward
    through the CUDA kernels: every decoder parameter gradient against the reference's."""
    g = golden_grpo
    c = g["cfg"]
    enc, head, dec = build_parts(c, 12)
    theta = M.GRPOViTOMR(enc, head, dec, g["tf_sd"])
    theta.load_state_dict(g["theta_sd"])
    theta = theta.to(DEV).train()
    for m in theta.modules():
        if hasattr(m, "dropout_p"):
            m.dropout_p = 0.0
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
        if isinstance(getattr(m, "dropout", None), float):                  # attention-probability dropout (MultiheadAttention)
            m.dropout = 0.0
    logits = theta.decoder(g["shifted"].to(DEV), g["xlatent"].to(DEV), g["attn_mask"].to(DEV), g["xmask"].to(DEV))
    valid = ~g["attn_mask"]
    assert torch.allclose(logits.detach().cpu()[valid], g["theta_logits"][valid], atol=1e-4, rtol=1e-3)
    obj = grpo.calc_grpo_objective(logits, g["rollouts"].to(DEV), g["attn_mask"].to(DEV), g["log_probs"].to(DEV),
                                   g["advantages"].to(DEV), c["epsilon"], c["num_groups"])
    bonus = grpo.calc_entropy_bonus(logits, g["attn_mask"].to(DEV), theta.decoder.vocab_size)
    assert abs(float(obj.detach()) - float(g["objective"])) < 1e-4 * max(1.0, abs(float(g["objective"])))
    assert abs(float(bonus.detach()) - float(g["bonus"])) < 1e-5
    (obj + c["entropy_beta"] * bonus).backward()
    bad = []
    for k, p in theta.decoder.named_parameters():
        ref = g["dec_grads"].get(k)
        if ref is None:
            continue
        got = p.grad.detach().float().cpu()
        if not torch.allclose(got, ref, atol=3e-5, rtol=2e-3):
            bad.append((k, float((got - ref).abs().max()), float(ref.abs().max())))
    assert not bad, bad[:5]


def test_engine_rollouts_match_module_path_and_oracle():
    """head_dim 64 decoder: the CUDA-graphed engine's rollouts (device-side draw inside the captured step) against the
    oracle's rollout on the same weights, fp32, explicit uniforms; then a seeded run replays bit for bit."""
    torch.manual_seed(2)
    dec = M.OMRDecoder(24, VOCAB, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, use_caching=True, max_batch_size=6,
                       cache_dtype=torch.float)
    enc = M.OMREncoder(16, 12, 40, num_layers=1, hidden_dim=128, num_heads=2, mlp_dim=256)
    vit = M.ViTOMR(enc, torch.nn.Identity(), dec)
    policy = M.GRPOViTOMR.__new__(M.GRPOViTOMR)
    M.ViTOMR.__init__(policy, enc, torch.nn.Identity(), dec)
    policy = policy.to(DEV).eval()
    g = torch.Generator().manual_seed(4)
    memory = torch.randn(6, 9, 128, generator=g)
    mem_mask = torch.zeros(6, 9, dtype=torch.bool)
    mem_mask[1, 6:] = True
    mem_mask[4, 3:] = True
    u = torch.rand(6, 16, generator=g)
    with torch.no_grad():
        rollouts, logp, rmask = policy.cached_forward_rollout_policy(memory.to(DEV), mem_mask.to(DEV), max_actions=16, top_k=20,
                                                                    temperature=1.1, uniforms=u)
    assert getattr(policy, "_acai_engines", None), "the decode engine should have taken this shape"
    sd = {k: v.detach().cpu() for k, v in dec.state_dict().items()}
    ro, lo, mo = O.cached_forward_rollout_policy(memory, sd, 2, u, mem_mask, max_actions=16, top_k=20, temperature=1.1,
                                                 pad_idx=dec.pad_idx, bos_idx=dec.bos_idx, eos_idx=dec.eos_idx)
    assert torch.equal(rollouts.cpu(), ro) and torch.equal(rmask.cpu(), mo)
    assert torch.allclose(logp.cpu(), lo, atol=1e-4, rtol=1e-3)
    with torch.no_grad():
        a = policy.cached_forward_rollout_policy(memory.to(DEV), mem_mask.to(DEV), max_actions=16, top_k=20, temperature=1.1, seed=7)
        b = policy.cached_forward_rollout_policy(memory.to(DEV), mem_mask.to(DEV), max_actions=16, top_k=20, temperature=1.1, seed=7)
        c2 = policy.cached_forward_rollout_policy(memory.to(DEV), mem_mask.to(DEV), max_actions=16, top_k=20, temperature=1.1, seed=8)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert a[0].shape != c2[0].shape or not torch.equal(a[0], c2[0])
