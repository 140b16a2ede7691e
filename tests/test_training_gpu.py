"""Backward kernels and the training steps (teacher-forced ViTOMR + OMRCELoss, MAE + MAELoss) against
(1) torch autograd on the same inputs (kernel level), (2) the reference's own parameter gradients stored in
tests/golden/golden_grads.pt (model level, fp32), (3) the CPU oracle's autograd at a mid size in bf16 mode."""
import math
import os

import pytest
import torch

from acai_omr_b200 import functional as F
from acai_omr_b200 import ops
from acai_omr_b200.layers import Packed
from acai_omr_b200.models import models as M
from oracle import vitomr_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
VOCAB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "acai_omr_b200", "data", "lmx_vocab.txt")


def rel_err(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-12))


def rnd(*shape, seed=0, dtype=torch.float32, scale=1.0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return (scale * torch.randn(*shape, device=DEV, generator=g)).to(dtype)


# ------------------------------------------------------------------------------------------ kernels
@pytest.mark.parametrize("shape", [(1, 1), (70, 33), (257, 768), (1000, 64), (5003, 3072), (513, 8)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_transpose_pad_and_colsum(shape, dtype):
    x = rnd(*shape, seed=1, dtype=dtype)
    t = ops.transpose_pad(x)
    R, Cc = shape
    Rp = (R + 63) // 64 * 64
    assert t.shape == (Cc, Rp)
    assert torch.equal(t[:, :R], x.t())
    assert not bool(t[:, R:].any())
    assert torch.equal(ops.transpose_pad(x, pad_to=1), x.t().contiguous())
    s = ops.colsum(x)
    assert torch.allclose(s, x.double().sum(0).float(), atol=1e-3 * math.sqrt(R), rtol=1e-4)
    s2 = ops.colsum(x)
    assert torch.equal(s, s2)          # fixed summation order
    acc = torch.ones(Cc, device=DEV)
    ops.colsum(x, out=acc, accumulate=True)
    assert torch.allclose(acc, 1.0 + s, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-6), (torch.bfloat16, 1e-2)])
def test_gelu_fwd_bwd(dtype, tol):
    z = rnd(333, 96, seed=2, dtype=dtype, scale=2.0)
    dy = rnd(333, 96, seed=3, dtype=dtype)
    zr = z.float().requires_grad_(True)
    yr = torch.nn.functional.gelu(zr)
    yr.backward(dy.float())
    assert rel_err(ops.gelu(z), yr) < tol
    assert rel_err(ops.gelu_bwd(dy, z), zr.grad) < tol


@pytest.mark.parametrize("rows,cols", [(5, 32), (300, 48), (1000, 768), (513, 1024), (64, 512)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("with_res", [False, True])
def test_layernorm_bwd(rows, cols, dtype, with_res):
    x = rnd(rows, cols, seed=4, dtype=dtype)
    res = rnd(rows, cols, seed=5, dtype=dtype) if with_res else None
    gamma, beta = 1.0 + 0.1 * rnd(cols, seed=6), 0.1 * rnd(cols, seed=7)
    dy = rnd(rows, cols, seed=8)
    v = (x.float() + (res.float() if with_res else 0.0)).requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    torch.nn.functional.layer_norm(v, (cols,), gr, br, 1e-6).backward(dy)
    dv, dg, db = ops.layernorm_bwd(dy, x, res, gamma, 1e-6)
    tol = 2e-5 if dtype == torch.float32 else 1e-2
    assert rel_err(dv, v.grad) < tol
    assert rel_err(dg, gr.grad) < tol and rel_err(db, br.grad) < 1e-5
    dv2, dg2, _ = ops.layernorm_bwd(dy, x, res, gamma, 1e-6)
    assert torch.equal(dg, dg2) and torch.equal(dv, dv2)


def torch_attention(q, k, v, lens_q, lens_k, H, causal):
    """fp32 reference over packed rows: per-sequence softmax(q k^T / sqrt(d)) v."""
    outs, oq, ok = [], 0, 0
    d = q.shape[1] // H
    for lq, lk in zip(lens_q, lens_k):
        qs = q[oq:oq + lq].view(lq, H, d).transpose(0, 1)
        ks = k[ok:ok + lk].view(lk, H, d).transpose(0, 1)
        vs = v[ok:ok + lk].view(lk, H, d).transpose(0, 1)
        s = qs @ ks.transpose(1, 2) / math.sqrt(d)
        if causal:
            s = s.masked_fill(torch.ones(lq, lk, dtype=torch.bool, device=s.device).triu(1), float("-inf"))
        outs.append((torch.softmax(s, -1) @ vs).transpose(0, 1).reshape(lq, H * d))
        oq, ok = oq + lq, ok + lk
    return torch.cat(outs)


@pytest.mark.parametrize("H,d", [(4, 8), (3, 12), (2, 32), (16, 32), (2, 64), (12, 64), (1, 128)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("causal", [False, True])
def test_self_attention_backward(H, d, dtype, causal):
    lens = [1, 70, 129, 33, 200]
    pk = Packed(lens, DEV)
    D = H * d
    qkv = rnd(pk.total, 3 * D, seed=9, dtype=dtype, scale=0.7).requires_grad_(True)
    dout = rnd(pk.total, D, seed=10, dtype=dtype)
    out = F.self_attention(qkv, pk, H, d, causal)
    out.backward(dout)
    ref_in = qkv.detach().float().requires_grad_(True)
    ref = torch_attention(ref_in[:, :D], ref_in[:, D:2 * D], ref_in[:, 2 * D:], lens, lens, H, causal)
    ref.backward(dout.float())
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    assert rel_err(out, ref) < tol
    for i, name in enumerate("qkv"):
        assert rel_err(qkv.grad[:, i * D:(i + 1) * D], ref_in.grad[:, i * D:(i + 1) * D]) < tol, name


@pytest.mark.parametrize("causal,p_drop", [(False, 0.0), (True, 0.0), (False, 0.1), (True, 0.1)])
def test_grouped_backward_kernels_match_two_cta_kernels_on_long_sequences(monkeypatch, causal, p_drop):
    """head_dim 64: the grouped dK/dV and dQ kernels (attn_bwd_g4: several trips round their 12 / 16-stage rings, ragged tails,
    causal first sub-blocks, dropout) against the two-CTA kernels on the same inputs, and both against fp32 torch math on the
    shortest sequence.  ACAI_ATTN_BWD_G4 selects the kernels per call."""
    H, d = 2, 64
    lens = [1031, 517, 2053, 64, 1]
    pk = Packed(lens, DEV)
    D = H * d
    qkv = rnd(pk.total, 3 * D, seed=31, dtype=torch.bfloat16, scale=0.7)
    dout = rnd(pk.total, D, seed=32, dtype=torch.bfloat16)
    kw = dict(dropout_p=p_drop, dropout_seed=77) if p_drop else {}
    grads = {}
    for flag in ("0", "1", "3"):
        monkeypatch.setenv("ACAI_ATTN_BWD_G4", flag)
        x = qkv.clone().requires_grad_(True)
        F.self_attention(x, pk, H, d, causal, **kw).backward(dout)
        grads[flag] = x.grad.float()
    for flag in ("1", "3"):
        for i, name in enumerate("qkv"):
            a, b = grads[flag][:, i * D:(i + 1) * D], grads["0"][:, i * D:(i + 1) * D]
            assert rel_err(a, b) < 4e-3, (flag, name, rel_err(a, b))          # same bf16 operands, fp32 sums in another order
    assert not torch.equal(grads["3"][:, :D], torch.zeros_like(grads["3"][:, :D]))
    if not p_drop:                                                             # torch math on one mid-sized sequence
        o0 = sum(lens[:1])
        sl = slice(o0, o0 + lens[1])
        ref_in = qkv[sl].float().requires_grad_(True)
        ref = torch_attention(ref_in[:, :D], ref_in[:, D:2 * D], ref_in[:, 2 * D:], [lens[1]], [lens[1]], H, causal)
        ref.backward(dout[sl].float())
        for i, name in enumerate("qkv"):
            assert rel_err(grads["3"][sl, i * D:(i + 1) * D], ref_in.grad[:, i * D:(i + 1) * D]) < 2e-2, name


@pytest.mark.parametrize("H,d", [(4, 12), (3, 32), (2, 64), (16, 64)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_cross_attention_backward(H, d, dtype):
    lens_q, lens_k = [5, 64, 1, 100], [130, 7, 300, 64]
    pq, pk = Packed(lens_q, DEV), Packed(lens_k, DEV)
    D = H * d
    q = rnd(pq.total, D, seed=11, dtype=dtype, scale=0.7).requires_grad_(True)
    kv = rnd(pk.total, 2 * D, seed=12, dtype=dtype, scale=0.7).requires_grad_(True)
    dout = rnd(pq.total, D, seed=13, dtype=dtype)
    out = F.cross_attention(q, kv, pq, pk, H, d)
    out.backward(dout)
    qr, kvr = q.detach().float().requires_grad_(True), kv.detach().float().requires_grad_(True)
    ref = torch_attention(qr, kvr[:, :D], kvr[:, D:], lens_q, lens_k, H, False)
    ref.backward(dout.float())
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    assert rel_err(out, ref) < tol
    assert rel_err(q.grad, qr.grad) < tol and rel_err(kv.grad, kvr.grad) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
def test_ce_loss_fwd_bwd(dtype, tol):
    B, T, V, pad = 5, 37, 227, 1
    logits = rnd(B, T, V, seed=14, dtype=dtype, scale=3.0).requires_grad_(True)
    g = torch.Generator().manual_seed(15)
    tgt = torch.randint(0, V, (B, T), generator=g)
    tgt[0, 20:] = pad
    tgt[3] = pad
    tgt = tgt.to(DEV)
    loss, sums = F.ce_loss(logits, tgt, pad)
    (loss * 1.7).backward()
    lr = logits.detach().float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lr.reshape(-1, V), tgt.reshape(-1), ignore_index=pad)
    (ref * 1.7).backward()
    assert abs(float(loss.detach()) - float(ref.detach())) < tol * abs(float(ref.detach()))
    assert float(sums[1]) == float((tgt != pad).sum())
    assert rel_err(logits.grad, lr.grad) < tol
    assert not bool(logits.grad[3].any())
    loss_g, _ = F.ce_loss(logits.detach(), tgt, pad, normaliser=torch.tensor([10.0], device=DEV))
    assert abs(float(loss_g) - float(sums[0]) / 10.0) < 1e-5 * abs(float(loss_g))


@pytest.mark.parametrize("eps", [0.1, 0.3])
def test_ce_loss_label_smoothing(eps):
    """OMRCELoss(pad_idx, label_smoothing) = nn.CrossEntropyLoss(ignore_index, label_smoothing) (models.py:785-788)."""
    from acai_omr_b200.models import models as M
    B, T, V, pad = 3, 17, 227, 1
    logits = rnd(B, T, V, seed=21, scale=2.0).requires_grad_(True)
    g = torch.Generator().manual_seed(22)
    tgt = torch.randint(0, V, (B, T), generator=g)
    tgt[1, 9:] = pad
    tgt = tgt.to(DEV)
    loss = M.OMRCELoss(pad, label_smoothing=eps)(logits, tgt)
    (loss * 0.7).backward()
    lr = logits.detach().clone().requires_grad_(True)
    ref = torch.nn.CrossEntropyLoss(ignore_index=pad, label_smoothing=eps)(lr.reshape(-1, V), tgt.reshape(-1))
    (ref * 0.7).backward()
    assert abs(float(loss.detach()) - float(ref.detach())) < 1e-5 * abs(float(ref.detach()))
    assert torch.allclose(logits.grad, lr.grad, atol=1e-7, rtol=1e-4)
    with torch.no_grad():                                   # the no-autograd path
        assert abs(float(M.OMRCELoss(pad, eps)(logits.detach(), tgt)) - float(ref.detach())) < 1e-5 * abs(float(ref.detach()))


def test_scatter_gather_rows():
    src = rnd(300, 48, seed=16)
    g = torch.Generator().manual_seed(17)
    idx = torch.randint(-1, 20, (300,), generator=g).to(DEV)
    dst = ops.scatter_add_rows(src, idx, torch.zeros(20, 48, device=DEV))
    ref = torch.zeros(21, 48, device=DEV).index_add_(0, torch.where(idx < 0, torch.full_like(idx, 20), idx), src)[:20]
    assert torch.allclose(dst, ref, atol=1e-4, rtol=1e-5)
    got = ops.gather_rows(src, idx)
    assert torch.equal(got[idx >= 0], src[idx[idx >= 0]]) and not bool(got[idx < 0].any())


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
def test_mae_loss_backward(dtype, tol):
    pred = rnd(3, 40, 256, seed=18, dtype=dtype).requires_grad_(True)
    target = torch.rand(3, 40, 256, device=DEV)
    mask = (torch.rand(3, 40, device=DEV) > 0.3)
    loss, _ = F.mae_loss(pred, mask, target)
    loss.backward()
    pr = pred.detach().float().cpu().requires_grad_(True)
    ref = O.mae_loss(pr, mask.cpu(), target.cpu())
    ref.backward()
    assert abs(float(loss.detach()) - float(ref.detach())) < tol * abs(float(ref.detach()))
    assert rel_err(pred.grad, pr.grad) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 1.5e-2)])
@pytest.mark.parametrize("M_,N,K,rows,act,res", [
    (300, 96, 64, None, 0, False), (129, 768, 768, None, 0, True), (500, 3072, 768, None, 1, False),
    (77, 227, 128, None, 0, False), (260, 128, 128, (0, 128), 0, False), (260, 256, 128, (128, 384), 0, False)])
def test_linear_backward(dtype, tol, M_, N, K, rows, act, res):
    Nw = N if rows is None else 384
    w = (rnd(Nw, K, seed=19) / math.sqrt(K)).requires_grad_(True)
    b = (0.1 * rnd(Nw, seed=20)).requires_grad_(True)
    x = rnd(M_, K, seed=21, dtype=dtype).requires_grad_(True)
    r = rnd(M_, N, seed=22, dtype=dtype).requires_grad_(True) if res else None
    dy = rnd(M_, N, seed=23, dtype=dtype)
    y = F.linear(x, w, b, act=act, residual=r, rows=rows)
    y.backward(dy)
    xr, wr, br = x.detach().float().requires_grad_(True), w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    ws, bs = (wr, br) if rows is None else (wr[rows[0]:rows[1]], br[rows[0]:rows[1]])
    if dtype == torch.bfloat16:
        ws = ws.detach().bfloat16().float() + (ws - ws.detach())        # forward with the bf16-rounded weight, gradient to the master
    yr = xr @ ws.t() + bs
    if act:
        yr = torch.nn.functional.gelu(yr)
    rr = None
    if res:
        rr = r.detach().float().requires_grad_(True)
        yr = yr + rr
    yr.backward(dy.float())
    assert rel_err(y, yr) < tol
    assert rel_err(x.grad, xr.grad) < tol and rel_err(w.grad, wr.grad) < tol and rel_err(b.grad, br.grad) < tol
    if res:
        assert rel_err(r.grad, rr.grad) < tol


# ------------------------------------------------------------------------------------------ model level
def cuda_list(ts):
    return [t.to(DEV) for t in ts]


def check_param_grads(module, ref_grads, atol, rtol, min_cos=None):
    bad = []
    for name, p in module.named_parameters():
        ref = ref_grads[name]
        assert p.grad is not None, name
        got = p.grad.detach().float().cpu()
        if min_cos is not None:
            if float(ref.norm()) < 1e-7:
                continue
            cos = float((got * ref).sum() / (got.norm() * ref.norm() + 1e-30))
            nr = float(got.norm() / ref.norm())
            if cos < min_cos or abs(nr - 1.0) > 1.0 - min_cos + 0.03:
                bad.append((name, cos, nr))
        elif not torch.allclose(got, ref, atol=atol, rtol=rtol):
            bad.append((name, float((got - ref).abs().max()), float(ref.abs().max())))
    assert not bad, bad


def test_teacher_forced_step_gradients_golden(golden, golden_grads):
    """One fp32 training step of TeacherForcedViTOMR + OMRCELoss: every parameter gradient vs the reference's."""
    g, gg = golden["vitomr"], golden_grads["vitomr"]
    c = g["cfg"]
    enc = M.FineTuneOMREncoder(4, 6, 10, 2, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64, transformer_dropout=0.0)
    dec = M.OMRDecoder(20, VOCAB, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96, transformer_dropout=0.0)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=c["transition_head_dim"], transition_head_dropout=0.0)
    vit.load_state_dict(g["sd"])
    for p in vit.parameters():
        p.requires_grad_(True)
    vit = vit.to(DEV).train()
    pred, tgt = vit(list(zip(cuda_list(g["imgs"]), cuda_list(g["lmx"]))))
    loss = M.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
    assert torch.allclose(loss.detach().cpu(), gg["loss"], atol=2e-5, rtol=1e-4)
    loss.backward()
    check_param_grads(vit, gg["grads"], atol=3e-5, rtol=2e-3)


def test_mae_step_gradients_golden(golden, golden_grads):
    g, gg = golden["mae"], golden_grads["mae"]
    c = g["cfg"]
    mae = M.MAE(c["mask_ratio"], c["patch_size"], c["pe_max_height"], c["pe_max_width"], c["encoder_hidden_dim"],
                c["decoder_hidden_dim"], c["encoder_kwargs"], c["decoder_kwargs"])
    mae.load_state_dict(g["sd"])
    mae = mae.to(DEV).train()
    pred, lmask, target = mae(list(zip(cuda_list(g["imgs"]), cuda_list(g["targets"]))), noises=g["noises"])
    loss = M.MAELoss()(pred, lmask, target)
    assert torch.allclose(loss.detach().cpu(), gg["loss"], atol=2e-5, rtol=1e-4)
    loss.backward()
    check_param_grads(mae, gg["grads"], atol=3e-5, rtol=2e-3)


def test_frozen_encoder_gets_no_gradients(golden):
    """The fine-tune recipe (models.py:663-677): frozen blocks / projection / PE stay without .grad."""
    enc = M.FineTuneOMREncoder(16, 12, 40, 1, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.0)
    dec = M.OMRDecoder(32, VOCAB, num_layers=1, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.0)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=128, transition_head_dropout=0.0).to(DEV).train()
    g = torch.Generator().manual_seed(3)
    imgs = [torch.rand(1, 64, 160, generator=g), torch.rand(1, 32, 320, generator=g)]
    lmx = [torch.tensor([0, 5, 9, 33, 2]), torch.tensor([0, 7, 2])]
    with torch.autocast("cuda", dtype=torch.bfloat16):
        pred, tgt = vit(list(zip(cuda_list(imgs), cuda_list(lmx))))
        loss = M.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
    loss.backward()
    for name, p in vit.named_parameters():
        frozen = name.startswith("encoder.frozen_blocks") or name.startswith("encoder.projection") or name == "encoder.pos_embedding"
        assert (p.grad is None) == frozen, name
        if p.grad is not None:
            assert p.grad.dtype == torch.float32 and bool(torch.isfinite(p.grad).all()), name


def _oracle_grads(loss_fn, sd):
    leaf = {k: v.clone().requires_grad_(True) if v.is_floating_point() else v for k, v in sd.items()}
    loss = loss_fn(leaf)
    loss.backward()
    return float(loss.detach()), {k: v.grad for k, v in leaf.items() if v.is_floating_point()}


def test_bf16_teacher_forced_step_vs_oracle_autograd():
    """Mid-size model with the real head dim (64 -> tcgen05 attention forward), whole step under bf16 autocast as
    omr_teacher_force_train.py runs it, against the fp32 CPU oracle differentiated by torch autograd."""
    from acai_omr_b200 import synth
    enc = M.FineTuneOMREncoder(16, 12, 40, 2, num_layers=3, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.0)
    dec = M.OMRDecoder(48, VOCAB, num_layers=2, hidden_dim=192, num_heads=3, mlp_dim=384, transformer_dropout=0.0)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=160, transition_head_dropout=0.0)
    synth.rerandomise_(vit, 41)
    for p in vit.parameters():
        p.requires_grad_(True)
    sd = {k: v.detach().clone() for k, v in vit.state_dict().items()}
    g = torch.Generator().manual_seed(42)
    imgs = [torch.rand(1, h, w, generator=g) for h, w in [(64, 160), (32, 320), (96, 96), (16, 640)]]
    lmx = [torch.cat([torch.tensor([0]), torch.randint(3, 227, (n,), generator=g), torch.tensor([2])]) for n in (30, 7, 45, 18)]
    enc_cfg = dict(patch_size=16, pe_max_height=12, pe_max_width=40, num_heads=2, interpolate=True,
                   stacks=["frozen_blocks", "fine_tune_blocks"])
    cfg = dict(encoder=enc_cfg, decoder_heads=3, pad_idx=1)

    def oracle_loss(leaf):
        pred, tgt = O.teacher_forced_forward(list(zip(imgs, lmx)), leaf, cfg)
        return O.ce_loss(pred, tgt, 1)

    loss_o, grads_o = _oracle_grads(oracle_loss, sd)
    vit = vit.to(DEV).train()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        pred, tgt = vit(list(zip(cuda_list(imgs), cuda_list(lmx))))
        loss = M.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
    loss.backward()
    assert abs(float(loss) - loss_o) < 2e-2 * abs(loss_o)
    check_param_grads(vit, grads_o, 0, 0, min_cos=0.97)


def test_bf16_mae_step_vs_oracle_autograd():
    from acai_omr_b200 import synth
    mae = M.MAE(0.75, 16, 12, 40, encoder_hidden_dim=128, decoder_hidden_dim=64,
                encoder_kwargs=dict(num_layers=2, num_heads=2, mlp_dim=256), decoder_kwargs=dict(num_layers=2, num_heads=2, mlp_dim=128))
    synth.rerandomise_(mae, 43)
    sd = {k: v.detach().clone() for k, v in mae.state_dict().items()}
    g = torch.Generator().manual_seed(44)
    imgs = [torch.rand(1, 64, 160, generator=g), torch.rand(1, 32, 320, generator=g), torch.rand(1, 48, 48, generator=g)]
    noises = [torch.rand((t.shape[1] // 16) * (t.shape[2] // 16), generator=g) for t in imgs]
    cfg = dict(mask_ratio=0.75, patch_size=16, pe_max_height=12, pe_max_width=40, encoder_heads=2, decoder_heads=2)

    def oracle_loss(leaf):
        return O.mae_loss(*O.mae_forward(list(zip(imgs, imgs)), leaf, cfg, noises))

    loss_o, grads_o = _oracle_grads(oracle_loss, sd)
    mae = mae.to(DEV).train()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        cu = cuda_list(imgs)
        loss = M.MAELoss()(*mae(list(zip(cu, cu)), noises=noises))
    loss.backward()
    assert abs(float(loss) - loss_o) < 2e-2 * abs(loss_o)
    check_param_grads(mae, grads_o, 0, 0, min_cos=0.97)


def test_adamw_kernel_matches_torch():
    from acai_omr_b200 import train
    torch.manual_seed(0)
    lin = torch.nn.Linear(300, 17).to(DEV)
    ref = torch.nn.Linear(300, 17).to(DEV)
    ref.load_state_dict(lin.state_dict())
    opt = torch.optim.AdamW(ref.parameters(), lr=3e-3, betas=(0.9, 0.95), weight_decay=0.01)
    tr = train.FlatTrainer(lin, lr=3e-3, betas=(0.9, 0.95), weight_decay=0.01)
    for it in range(4):
        gs = [rnd(*p.shape, seed=50 + it + i) for i, p in enumerate(lin.parameters())]
        tr.zero_grad()
        for p, r, g in zip(lin.parameters(), ref.parameters(), gs):
            p.grad.copy_(g)
            r.grad = g.clone()
        tr.step()
        opt.step()
    for p, r in zip(lin.parameters(), ref.parameters()):
        assert torch.allclose(p, r, atol=1e-6, rtol=1e-5)
    assert torch.equal(tr.shadow[: lin.weight.numel()].view_as(lin.weight), lin.weight.detach().bfloat16())


def test_trainer_steps_reduce_the_loss_and_refresh_weight_copies():
    """Three fused steps on a fixed batch: loss falls, the bf16 shadow / transposed copies follow the parameters."""
    from acai_omr_b200 import layers, synth, train
    enc = M.FineTuneOMREncoder(16, 12, 40, 1, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.0)
    dec = M.OMRDecoder(48, VOCAB, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.0)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=128, transition_head_dropout=0.0)
    synth.rerandomise_(vit, 5)
    vit = vit.to(DEV).train()
    groups, lrs = vit.create_fine_tune_param_groups(2e-3, 1e-3, 0.9)
    tr = train.FlatTrainer(vit, groups)
    frozen_before = vit.encoder.frozen_blocks.layers[0].linear1.weight.detach().clone()
    imgs = [synth.synth_system(h, w, i).to(DEV) for i, (h, w) in enumerate([(64, 160), (32, 320), (96, 96)])]
    batch = list(zip(imgs, [synth.synth_lmx(L, i).to(DEV) for i, L in enumerate([20, 33, 12])]))
    losses = []
    for _ in range(4):
        sums = train.teacher_forced_step(vit, batch, tr)
        losses.append(float(sums[0] / sums[1]))
    assert losses[-1] < losses[0] - 0.05, losses
    w = vit.decoder.unembed.weight
    assert torch.equal(layers.w_(w, torch.bfloat16), w.detach().bfloat16())
    assert torch.equal(layers.WEIGHTS.get_t(w, torch.bfloat16), w.detach().bfloat16().t().contiguous())
    assert torch.equal(vit.encoder.frozen_blocks.layers[0].linear1.weight, frozen_before)
    layers.WEIGHTS.clear()


def test_trainer_checkpoint_resume():
    """the run is resumable - model state_dict (reference key names) + FlatTrainer.state_dict (AdamW moments,
    step, per-group learning rates) restored into fresh objects continue exactly like the uninterrupted run (up to the fp32
    reduce-add order of the split-K weight gradients)."""
    from acai_omr_b200 import layers, synth, train

    def make():
        enc = M.FineTuneOMREncoder(16, 12, 40, 1, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.0)
        dec = M.OMRDecoder(48, VOCAB, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.0)
        vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=128, transition_head_dropout=0.0)
        synth.rerandomise_(vit, 6)
        vit = vit.to(DEV).train()
        groups, _ = vit.create_fine_tune_param_groups(2e-3, 1e-3, 0.9)
        return vit, train.FlatTrainer(vit, groups)

    imgs = [synth.synth_system(h, w, i).to(DEV) for i, (h, w) in enumerate([(64, 160), (32, 320)])]
    batch = list(zip(imgs, [synth.synth_lmx(L, i).to(DEV) for i, L in enumerate([20, 33])]))
    loss = lambda sums: float(sums[0] / sums[1])  # noqa: E731
    vit, tr = make()
    for _ in range(2):
        train.teacher_forced_step(vit, batch, tr)
    ckpt = {"model": {k: v.detach().clone() for k, v in vit.state_dict().items()}, "trainer": tr.state_dict()}
    cont = [loss(train.teacher_forced_step(vit, batch, tr)) for _ in range(3)]
    layers.WEIGHTS.clear()
    vit2, tr2 = make()
    vit2.load_state_dict(ckpt["model"])
    tr2.load_state_dict(ckpt["trainer"])
    assert tr2.step_count == 2
    resumed = [loss(train.teacher_forced_step(vit2, batch, tr2)) for _ in range(3)]
    assert torch.allclose(torch.tensor(resumed), torch.tensor(cont), rtol=2e-3, atol=2e-3), (cont, resumed)
    assert resumed[-1] < resumed[0]
    layers.WEIGHTS.clear()


@pytest.mark.parametrize("T,n_out,k_in", [(1000, 768, 768), (63, 128, 256), (5000, 3072, 768), (777, 200, 136), (4096, 64, 64),
                                          (20000, 1024, 4096), (130, 8, 1024)])
def test_wgrad_accumulate_tensor_core(T, n_out, k_in):
    """dW += dZ^T X straight from token-major operands (MN-major UMMA operands, split-K, TMA reduce-add)."""
    dz = rnd(T, n_out, seed=60, dtype=torch.bfloat16)
    x = rnd(T, k_in, seed=61, dtype=torch.bfloat16)
    dw0 = rnd(n_out, k_in, seed=62)
    dw = dw0.clone()
    assert ops.wgrad_supported(dz, x, dw)
    ops.wgrad_accumulate(dz, x, dw)
    ref = dw0.double() + dz.double().t() @ x.double()
    # (a token range is one fp32 accumulation chain on the tensor core, whose adds truncate: 1.1e-5 per 10 k tokens against
    # fp64, chains are capped at 32 k tokens - profiles/ubench/wgrad_precision.py; a bf16-output matmul of the same operands is
    # at 1.7e-3)
    assert rel_err(dw, ref) < 4e-5
    # strided views (a slice of a fused QKV gradient buffer) and a row slice of the destination
    big = rnd(T, 3 * n_out, seed=63, dtype=torch.bfloat16)
    if n_out % 8 == 0:
        dwf = torch.zeros(3 * n_out, k_in, device=DEV)
        ops.wgrad_accumulate(big[:, n_out:2 * n_out], x, dwf[n_out:2 * n_out])
        assert rel_err(dwf[n_out:2 * n_out], big[:, n_out:2 * n_out].double().t() @ x.double()) < 4e-5
        assert not bool(dwf[:n_out].any()) and not bool(dwf[2 * n_out:].any())


@pytest.mark.parametrize("T,n_out,k_in", [(1000, 768, 768), (63, 128, 256), (5000, 3072, 768), (777, 200, 136), (4096, 64, 64),
                                          (20000, 1024, 4096), (130, 8, 1024), (1, 264, 72), (64 * 37 + 1, 2304, 768)])
def test_wgrad_with_bias_gradient_from_the_same_pass(T, n_out, k_in):
    """acai_wgrad_bias_accumulate: dW as acai_wgrad_accumulate AND db += column sums of dZ, summed from the dZ boxes the
    pipeline stages in shared memory (ragged token / feature tails are TMA zero fill); against fp64 and acai_colsum."""
    dz = rnd(T, n_out, seed=64, dtype=torch.bfloat16)
    x = rnd(T, k_in, seed=65, dtype=torch.bfloat16)
    dw0, db0 = rnd(n_out, k_in, seed=66), rnd(n_out, seed=67)
    dw, db = dw0.clone(), db0.clone()
    assert ops.wgrad_supported(dz, x, dw) and ops.FUSE_BIAS_GRAD
    n0 = ops.launch_count()
    ops.wgrad_accumulate(dz, x, dw, dbias=db)
    assert ops.launch_count() - n0 == 1                                      # one kernel: no column-sum pass
    assert rel_err(dw, dw0.double() + dz.double().t() @ x.double()) < 4e-5
    assert rel_err(db, db0.double() + dz.double().sum(0)) < 1e-5
    assert rel_err(db - db0, ops.colsum(dz)) < 1e-5
    # a strided dZ view (the K slice of a fused QKV gradient) into row slices of the destinations
    if n_out % 8 == 0:
        big = rnd(T, 3 * n_out, seed=68, dtype=torch.bfloat16)
        dwf, dbf = torch.zeros(3 * n_out, k_in, device=DEV), torch.zeros(3 * n_out, device=DEV)
        ops.wgrad_accumulate(big[:, n_out:2 * n_out], x, dwf[n_out:2 * n_out], dbias=dbf[n_out:2 * n_out])
        assert rel_err(dbf[n_out:2 * n_out], big[:, n_out:2 * n_out].double().sum(0)) < 1e-5
        assert not bool(dbf[:n_out].any()) and not bool(dbf[2 * n_out:].any())
        assert rel_err(dwf[n_out:2 * n_out], big[:, n_out:2 * n_out].double().t() @ x.double()) < 4e-5


def test_linear_autograd_bias_gradient_fused_and_separate_agree(monkeypatch):
    """functional.linear's backward with the bias gradient inside the weight-gradient kernel (default) and through acai_colsum."""
    x = rnd(300, 256, seed=70, dtype=torch.bfloat16)
    w = rnd(520, 256, seed=71).requires_grad_(True)
    b = rnd(520, seed=72).requires_grad_(True)
    dy = rnd(300, 520, seed=73, dtype=torch.bfloat16)
    grads = []
    for fuse in (True, False):
        monkeypatch.setattr(ops, "FUSE_BIAS_GRAD", fuse)
        w.grad = b.grad = None
        F.linear(x, w, b, rows=(8, 520 - 8)).backward(dy[:, 8:-8])
        grads.append((w.grad.clone(), b.grad.clone()))
    assert rel_err(grads[0][0], grads[1][0]) < 1e-6 and rel_err(grads[0][1], grads[1][1]) < 1e-5
    assert not bool(grads[0][1][:8].any()) and not bool(grads[0][1][-8:].any())
    assert rel_err(grads[0][1][8:-8], dy[:, 8:-8].double().sum(0)) < 1e-5


@pytest.mark.parametrize("hard", [False, True])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 1.5e-2)])
def test_gumbel_softmax_fwd_bwd(hard, dtype, tol):
    logits = rnd(3, 11, 227, seed=80, dtype=dtype, scale=2.0).requires_grad_(True)
    noise = -torch.empty(3, 11, 227, device=DEV).exponential_(generator=torch.Generator(device=DEV).manual_seed(81)).log()
    dy = rnd(3, 11, 227, seed=82, dtype=dtype)
    y = F.gumbel_softmax(logits, noise, 1.5, hard)
    y.backward(dy)
    lr = logits.detach().float().requires_grad_(True)
    ref = O.gumbel_softmax(lr, noise, 1.5, hard)
    ref.backward(dy.float())
    assert rel_err(y, ref) < tol
    assert rel_err(logits.grad, lr.grad) < tol
    if hard:
        assert torch.equal(y.detach().float().sum(-1), torch.ones(3, 11, device=DEV)) and bool(((y == 0) | (y == 1)).all())


def test_mix_rows_and_expected_embedding_backward():
    B, T, D, V = 3, 7, 48, 227
    gold = rnd(B, T, D, seed=83).requires_grad_(True)
    exp = rnd(B, T, D, seed=84).requires_grad_(True)
    pos = rnd(20, D, seed=85).requires_grad_(True)
    mask = torch.rand(B, T, device=DEV, generator=torch.Generator(device=DEV).manual_seed(86)) < 0.5
    row = torch.arange(B * T, device=DEV).view(B, T)
    sel = torch.where(mask & (torch.arange(T, device=DEV) > 0)[None, :], row - 1, torch.full_like(row, -1))
    out = F.mix_rows(F.mix_rows(gold, a=exp, sel=sel), pos=pos, period=T)
    dy = rnd(B, T, D, seed=87)
    out.backward(dy)
    g2, e2, p2 = (t.detach().clone().requires_grad_(True) for t in (gold, exp, pos))
    shifted = torch.cat([g2[:, :1], e2], dim=1)[:, :-1]
    ref = torch.where(mask.unsqueeze(-1), shifted, g2) + p2[:T].unsqueeze(0)
    ref.backward(dy)
    assert torch.allclose(out, ref, atol=1e-6)
    for a, b in ((gold, g2), (exp, e2), (pos, p2)):
        assert torch.allclose(a.grad, b.grad, atol=1e-5), float((a.grad - b.grad).abs().max())
    # probs @ table and its two gradients
    probs = torch.softmax(rnd(B * T, V, seed=88), -1).requires_grad_(True)
    table = rnd(V, D, seed=89).requires_grad_(True)
    out = F.expected_embedding(probs, table)
    dy2 = rnd(B * T, D, seed=90)
    out.backward(dy2)
    pr, tr = probs.detach().clone().requires_grad_(True), table.detach().clone().requires_grad_(True)
    (pr @ tr).backward(dy2)
    assert rel_err(out, pr @ tr) < 1e-5 and rel_err(probs.grad, pr.grad) < 1e-5 and rel_err(table.grad, tr.grad) < 1e-5


@pytest.mark.parametrize("mode", ["soft", "hard"])
def test_scheduled_sampling_step_golden(golden, golden_grads, mode):
    """ScheduledSamplingViTOMR.forward_train + OMRCELoss + backward in fp32 with the reference's random draws replayed:
    predictions, loss and every parameter gradient vs the reference's."""
    g, gg = golden["vitomr"], golden_grads["scheduled_sampling"][mode]
    c = g["cfg"]
    enc = M.FineTuneOMREncoder(4, 6, 10, 2, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64, transformer_dropout=0.0)
    dec = M.OMRDecoder(20, VOCAB, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96, transformer_dropout=0.0)
    vit = M.ScheduledSamplingViTOMR(enc, None, dec, transition_head_dim=c["transition_head_dim"], transition_head_dropout=0.0)
    vit.load_state_dict(g["sd"])
    for p in vit.parameters():
        p.requires_grad_(True)
    vit = vit.to(DEV).train()
    pred, tgt = vit.forward_train(list(zip(cuda_list(g["imgs"]), cuda_list(g["lmx"]))), gg["tf_prob"], gg["tau"], mode == "hard",
                                  sample_rand=gg["rand"], gumbel_noise=gg["gumbel"])
    assert torch.equal(tgt.cpu(), gg["target"])
    valid = gg["target"] != 1
    assert torch.allclose(pred.detach().cpu()[valid], gg["pred"][valid], atol=3e-5, rtol=1e-3)
    loss = M.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
    assert torch.allclose(loss.detach().cpu(), gg["loss"], atol=2e-5, rtol=1e-4)
    loss.backward()
    check_param_grads(vit, gg["grads"], atol=3e-5, rtol=2e-3)
    with torch.no_grad():
        p_eval, _ = vit.forward_eval(list(zip(cuda_list(g["imgs"]), cuda_list(g["lmx"]))))
    assert p_eval.shape == pred.shape


# ------------------------------------------------------------------------------------------ dropout
def test_hidden_dropout_matches_exported_mask_and_rate():
    p, seed = 0.1, 0x1234_5678_9ABC_DEF0
    x = rnd(1000, 771, seed=91).requires_grad_(True)            # odd element count: last hash pair is half used
    res = rnd(1000, 771, seed=92).requires_grad_(True)
    keep = ops.dropout_mask(x.numel(), p, seed, DEV).view(x.shape)
    y = F._Dropout.apply(x, res, p, seed)
    dy = rnd(1000, 771, seed=93)
    y.backward(dy)
    assert torch.allclose(y, res + keep * x / (1 - p), atol=1e-6)
    assert torch.allclose(x.grad, keep * dy / (1 - p), atol=1e-6) and torch.equal(res.grad, dy)
    rate = 1.0 - keep.float().mean().item()
    assert abs(rate - p) < 4 * math.sqrt(p * (1 - p) / x.numel()) + 1e-5, rate        # 4 sigma + threshold quantisation
    other = ops.dropout_mask(x.numel(), p, ops.DropoutState.next_seed(), DEV).view(x.shape)
    agree = (other == keep).float().mean().item()
    assert abs(agree - (p * p + (1 - p) ** 2)) < 0.01                                # independent masks for different seeds
    # rows / columns are not correlated
    assert abs(keep.float().mean(0).std().item() - math.sqrt(p * (1 - p) / 1000)) < 0.003
    xb = rnd(64, 256, seed=94, dtype=torch.bfloat16)
    kb = ops.dropout_mask(xb.numel(), 0.05, 7, DEV).view(xb.shape)
    assert torch.equal(ops.dropout(xb, 0.05, 7), (kb * xb.float() / 0.95).bfloat16())


def masked_attention_reference(q, k, v, lens_q, lens_k, H, causal, keep, p):
    """fp32 torch attention with an explicit probability-dropout mask keep[h, q_global, k_local]."""
    outs, oq, ok = [], 0, 0
    d = q.shape[1] // H
    for lq, lk in zip(lens_q, lens_k):
        qs = q[oq:oq + lq].view(lq, H, d).transpose(0, 1)
        ks = k[ok:ok + lk].view(lk, H, d).transpose(0, 1)
        vs = v[ok:ok + lk].view(lk, H, d).transpose(0, 1)
        s = qs @ ks.transpose(1, 2) / math.sqrt(d)
        if causal:
            s = s.masked_fill(torch.ones(lq, lk, dtype=torch.bool, device=s.device).triu(1 + lk - lq), float("-inf"))
        pr = torch.softmax(s, -1) * keep[:, oq:oq + lq, :lk] / (1 - p)
        outs.append((pr @ vs).transpose(0, 1).reshape(lq, H * d))
        oq, ok = oq + lq, ok + lk
    return torch.cat(outs)


@pytest.mark.parametrize("H,d,dtype", [(4, 8, torch.float32), (3, 12, torch.float32), (2, 32, torch.bfloat16), (2, 64, torch.bfloat16),
                                       (16, 64, torch.bfloat16), (2, 64, torch.float32)])
@pytest.mark.parametrize("causal", [False, True])
def test_attention_probability_dropout_fwd_bwd(H, d, dtype, causal):
    """Forward and both backward passes regenerate the same mask (exported by acai_attn_dropout_mask): exact comparison
    against masked fp32 attention, tensor-core (bf16, d 32/64) and SIMT kernels alike."""
    p, seed = 0.1, 0xACA1_0000_0000_0042
    lens = [1, 70, 129, 33, 200]
    pk = Packed(lens, DEV)
    D = H * d
    qkv = rnd(pk.total, 3 * D, seed=95, dtype=dtype, scale=0.7).requires_grad_(True)
    dout = rnd(pk.total, D, seed=96, dtype=dtype)
    out = F.self_attention(qkv, pk, H, d, causal, dropout_p=p, dropout_seed=seed)
    out.backward(dout)
    keep = ops.attn_dropout_mask(p, seed, H, pk.total, pk.max_len, DEV).float()
    rate = 1.0 - keep.mean().item()
    assert abs(rate - p) < 0.005, rate
    ref_in = qkv.detach().float().requires_grad_(True)
    ref = masked_attention_reference(ref_in[:, :D], ref_in[:, D:2 * D], ref_in[:, 2 * D:], lens, lens, H, causal, keep, p)
    ref.backward(dout.float())
    tol = 3e-5 if dtype == torch.float32 else 2e-2
    assert rel_err(out, ref) < tol
    for i, name in enumerate("qkv"):
        assert rel_err(qkv.grad[:, i * D:(i + 1) * D], ref_in.grad[:, i * D:(i + 1) * D]) < tol, name


@pytest.mark.parametrize("H,d,dtype", [(4, 12, torch.float32), (3, 32, torch.bfloat16), (16, 64, torch.bfloat16)])
def test_cross_attention_probability_dropout(H, d, dtype):
    p, seed = 0.25, 99
    lens_q, lens_k = [5, 64, 1, 100], [130, 7, 300, 64]
    pq, pk = Packed(lens_q, DEV), Packed(lens_k, DEV)
    D = H * d
    q = rnd(pq.total, D, seed=97, dtype=dtype, scale=0.7).requires_grad_(True)
    kv = rnd(pk.total, 2 * D, seed=98, dtype=dtype, scale=0.7).requires_grad_(True)
    dout = rnd(pq.total, D, seed=99, dtype=dtype)
    out = F.cross_attention(q, kv, pq, pk, H, d, dropout_p=p, dropout_seed=seed)
    out.backward(dout)
    keep = ops.attn_dropout_mask(p, seed, H, pq.total, pk.max_len, DEV).float()
    qr, kvr = q.detach().float().requires_grad_(True), kv.detach().float().requires_grad_(True)
    ref = masked_attention_reference(qr, kvr[:, :D], kvr[:, D:], lens_q, lens_k, H, False, keep, p)
    ref.backward(dout.float())
    tol = 3e-5 if dtype == torch.float32 else 2e-2
    assert rel_err(out, ref) < tol and rel_err(q.grad, qr.grad) < tol and rel_err(kv.grad, kvr.grad) < tol


def test_training_mode_with_reference_dropout_rates():
    """The reference's recipe (encoder 0.05, transition head 0.05, decoder 0.1; omr_teacher_force_train.py:43-46) runs in
    train mode: finite loss and gradients, a different mask every step, eval mode unaffected and deterministic."""
    from acai_omr_b200 import synth
    ops.DropoutState.manual_seed(1)
    enc = M.FineTuneOMREncoder(16, 12, 40, 2, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.05)
    dec = M.OMRDecoder(48, VOCAB, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.1)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=128, transition_head_dropout=0.05)
    synth.rerandomise_(vit, 6)
    vit = vit.to(DEV).train()
    imgs = [synth.synth_system(h, w, i).to(DEV) for i, (h, w) in enumerate([(64, 160), (32, 320), (96, 96)])]
    batch = list(zip(imgs, [synth.synth_lmx(L, i).to(DEV) for i, L in enumerate([20, 33, 12])]))
    losses = []
    for _ in range(3):
        vit.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            pred, tgt = vit(batch)
            loss = M.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
        loss.backward()
        losses.append(float(loss.detach()))
        assert all(bool(torch.isfinite(p.grad).all()) for p in vit.parameters() if p.grad is not None)
    assert len(set(losses)) == 3, losses                      # fresh masks per step
    vit.eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        a, _ = vit(batch)
        b, _ = vit(batch)
    assert torch.equal(a, b)
    ops.DropoutState.manual_seed(1)
    vit.train()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        first = float(M.OMRCELoss(vit.decoder.pad_idx)(*vit(batch)).detach())
    assert first == losses[0]                                  # manual_seed replays the run


@pytest.mark.parametrize("autocast", [False, True])
def test_residual_gradient_hand_over_matches_autograd_adds(monkeypatch, autocast):
    """functional.GradBox: the residual gradient of every post-norm block (self-attention, cross-attention, feed-forward with
    its output dropout) is added inside the block's first dX GEMM instead of by an autograd add - same masks (replayed seed),
    same gradients as the plain autograd path up to the one bf16 rounding the fused add saves."""
    from acai_omr_b200 import synth
    from acai_omr_b200 import functional as Fn
    enc = M.FineTuneOMREncoder(16, 12, 40, 2, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.05)
    dec = M.OMRDecoder(48, VOCAB, num_layers=2, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.1)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=128, transition_head_dropout=0.05)
    synth.rerandomise_(vit, 6)
    vit = vit.to(DEV).train()
    imgs = [synth.synth_system(h, w, i).to(DEV) for i, (h, w) in enumerate([(64, 160), (32, 320), (96, 96)])]
    batch = list(zip(imgs, [synth.synth_lmx(L, i).to(DEV) for i, L in enumerate([20, 33, 12])]))
    runs = []
    for box in (True, False):
        monkeypatch.setattr(Fn, "GRAD_BOX", box)
        ops.DropoutState.manual_seed(3)
        vit.zero_grad()
        n0 = ops.launch_count()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            pred, tgt = vit(batch)
            loss = M.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
        loss.backward()
        runs.append((float(loss.detach()), {n: p.grad.clone() for n, p in vit.named_parameters() if p.grad is not None},
                     ops.launch_count() - n0))
    assert runs[0][0] == runs[1][0]                            # identical forward
    assert runs[0][1].keys() == runs[1][1].keys()
    tol = 2e-2 if autocast else 1e-5
    for name, g in runs[0][1].items():
        assert rel_err(g, runs[1][1][name]) < tol, name


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 1.5e-2)])
def test_gelu_linear_with_fused_dropout(dtype, tol):
    """Linear -> GELU -> Dropout as one autograd node (dropout fused into the GELU kernels, mask regenerated in the backward)."""
    ops.DropoutState.manual_seed(5)
    seed = ops.DropoutState.next_seed()
    ops.DropoutState.manual_seed(5)                      # F.linear will draw the same seed
    M_, N, K, p = 300, 256, 128, 0.2
    w = (rnd(N, K, seed=101) / math.sqrt(K)).requires_grad_(True)
    b = (0.1 * rnd(N, seed=102)).requires_grad_(True)
    x = rnd(M_, K, seed=103, dtype=dtype).requires_grad_(True)
    dy = rnd(M_, N, seed=104, dtype=dtype)
    y = F.linear(x, w, b, act=ops.ACT_GELU, dropout_p=p)
    y.backward(dy)
    keep = ops.dropout_mask(M_ * N, p, seed, DEV).view(M_, N).float()
    xr, wr, br = x.detach().float().requires_grad_(True), w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    ws = wr if dtype == torch.float32 else wr.detach().bfloat16().float() + (wr - wr.detach())
    z = xr @ ws.t() + br
    if dtype == torch.bfloat16:
        z = z.detach().bfloat16().float() + (z - z.detach())      # the kept pre-activation is bf16
    yr = torch.nn.functional.gelu(z) * keep / (1 - p)
    yr.backward(dy.float())
    assert rel_err(y, yr) < tol and rel_err(x.grad, xr.grad) < tol and rel_err(w.grad, wr.grad) < tol and rel_err(b.grad, br.grad) < tol


# ------------------------------------------------------------------------------------------ GELU in the GEMM epilogues
@pytest.mark.parametrize("M_,N,K", [(1000, 3072, 768), (129, 256, 64), (4100, 768, 3072), (37, 136, 128)])
@pytest.mark.parametrize("p", [0.0, 0.25])
def test_linear_gelu_train_epilogue(M_, N, K, p):
    """acai_linear_gelu_train: z is the plain GEMM's output bit for bit, h = dropout(gelu(z)) with the standalone kernel's mask."""
    x = rnd(M_, K, seed=111, dtype=torch.bfloat16)
    w = (rnd(N, K, seed=112) / math.sqrt(K)).bfloat16()
    b = 0.1 * rnd(N, seed=113)
    seed = 0x1234567887654321
    zh = ops.linear_gelu_train(x, w, b, p, seed)
    assert zh is not None, "the TMA-store epilogue must take this shape"
    z, h = zh
    z_ref = ops.linear(x, w, b, out_dtype=torch.bfloat16)
    assert torch.equal(z, z_ref)
    h_ref = ops.gelu(z_ref, dropout_p=p, dropout_seed=seed)
    if p > 0.0:
        keep = ops.dropout_mask(M_ * N, p, seed, DEV).view(M_, N)
        inner = (z.float().abs() < 4.0) & (z != 0)                 # gelu is only 0 at z = 0 or where 1 + erf underflows
        assert torch.equal((h != 0)[inner], keep[inner]) and float(h[~keep].abs().max()) == 0.0
    # Abramowitz-Stegun erf (1.5e-7) vs erff: at most the last bf16 bit
    assert float((h.float() - h_ref.float()).abs().max()) <= 2.0 ** -7 * float(h_ref.float().abs().max())
    assert rel_err(h, h_ref) < 2e-3


@pytest.mark.parametrize("M_,N,K", [(1000, 3072, 768), (129, 256, 64), (37, 136, 128)])
@pytest.mark.parametrize("p", [0.0, 0.25])
def test_linear_gelu_bwd_epilogue(M_, N, K, p):
    """acai_linear_gelu_bwd against fp32 torch: dz = (dy @ wt^T) * gelu'(z) * mask / (1 - p)."""
    dy = rnd(M_, K, seed=121, dtype=torch.bfloat16)
    wt = (rnd(N, K, seed=122) / math.sqrt(K)).bfloat16()
    z = rnd(M_, N, seed=123, dtype=torch.bfloat16, scale=1.5)
    seed = 0x0badcafe12345678
    dz = ops.linear_gelu_bwd(dy, wt, z, p, seed)
    assert dz is not None
    zf = z.float().requires_grad_(True)
    torch.nn.functional.gelu(zf).sum().backward()
    ref = (dy.float() @ wt.float().t()) * zf.grad
    if p > 0.0:
        keep = ops.dropout_mask(M_ * N, p, seed, DEV).view(M_, N)
        ref = ref * keep / (1 - p)
        assert float(dz[~keep].abs().max()) == 0.0
    assert rel_err(dz, ref) < 4e-3
    unfused = ops.gelu_bwd(ops.linear(dy, wt, out_dtype=torch.bfloat16), z, dropout_p=p, dropout_seed=seed)
    assert rel_err(dz, ref) <= rel_err(unfused, ref) * 1.05 + 1e-4       # fp32 accumulators into the derivative: no worse


@pytest.mark.parametrize("M_,N,K", [(1000, 768, 768), (513, 1024, 4096), (300, 128, 64)])
@pytest.mark.parametrize("with_res", [True, False])
def test_linear_dropout_epilogue(M_, N, K, with_res):
    """acai_linear_dropout: residual + dropout(x W^T + b) from one GEMM, the mask being acai_dropout's on the dense output."""
    x = rnd(M_, K, seed=131, dtype=torch.bfloat16)
    w = (rnd(N, K, seed=132) / math.sqrt(K)).bfloat16()
    b = 0.1 * rnd(N, seed=133)
    res = rnd(M_, N, seed=134, dtype=torch.bfloat16) if with_res else None
    p, seed = 0.1, 0x5eed0000beef
    y = ops.linear_dropout(x, w, b, res, p, seed)
    keep = ops.dropout_mask(M_ * N, p, seed, DEV).view(M_, N)
    ref = (x.float() @ w.float().t() + b) * keep / (1 - p) + (res.float() if with_res else 0.0)
    unfused = ops.dropout(ops.linear(x, w, b, out_dtype=torch.bfloat16), p, seed, residual=res)
    if y is not None:                    # (small problems are served by tiles without the TMA-store epilogue: not fused)
        assert rel_err(y, ref) < 4e-3
        assert rel_err(y, ref) <= rel_err(unfused, ref) * 1.05 + 1e-4
        if not with_res:
            assert float(y[~keep].abs().max()) == 0.0
    else:
        assert M_ * N < 148 * 128 * 128
    # through autograd (fused or not): the gradients of residual + dropout(linear)
    xr = x.clone().requires_grad_(True)
    wr = w.float().clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    ops.DropoutState.manual_seed(5)
    out = F.linear(xr, wr, br, residual=res, out_dtype=torch.bfloat16, dropout_p=p)
    dy = rnd(M_, N, seed=135, dtype=torch.bfloat16)
    out.backward(dy)
    ops.DropoutState.manual_seed(5)
    keep2 = ops.dropout_mask(M_ * N, p, ops.DropoutState.next_seed(), DEV).view(M_, N)
    assert rel_err(out, (x.float() @ w.float().t() + b) * keep2 / (1 - p) + (res.float() if with_res else 0.0)) < 4e-3
    dz = dy.float() * keep2 / (1 - p)
    assert rel_err(xr.grad, dz @ w.float()) < 1e-2
    assert rel_err(wr.grad, dz.t() @ x.float()) < 1e-2
    assert rel_err(br.grad, dz.sum(0)) < 1e-2


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 1.5e-2)])
@pytest.mark.parametrize("p,add_residual", [(0.0, True), (0.2, False)])
def test_feed_forward_node(dtype, tol, p, add_residual):
    """The feed-forward block as one tape node (fused epilogues in bf16, plain kernels in fp32) against torch autograd."""
    ops.DropoutState.manual_seed(9)
    seed = ops.DropoutState.next_seed()
    ops.DropoutState.manual_seed(9)
    T, d, hdim = 333, 128, 512
    w1 = (rnd(hdim, d, seed=131) / math.sqrt(d)).requires_grad_(True)
    b1 = (0.1 * rnd(hdim, seed=132)).requires_grad_(True)
    w2 = (rnd(d, hdim, seed=133) / math.sqrt(hdim)).requires_grad_(True)
    b2 = (0.1 * rnd(d, seed=134)).requires_grad_(True)
    x = rnd(T, d, seed=135, dtype=dtype).requires_grad_(True)
    dy = rnd(T, d, seed=136, dtype=dtype)
    y = F.feed_forward(x, w1, b1, w2, b2, add_residual=add_residual, dropout_p=p)
    y.backward(dy)
    rq = (lambda t: t) if dtype == torch.float32 else (lambda t: t.detach().bfloat16().float() + (t - t.detach()))
    xr = x.detach().float().requires_grad_(True)
    ps = [t.detach().clone().requires_grad_(True) for t in (w1, b1, w2, b2)]
    z = rq(xr @ rq(ps[0]).t() + ps[1])
    h = torch.nn.functional.gelu(z)
    if p > 0.0:
        h = h * ops.dropout_mask(T * hdim, p, seed, DEV).view(T, hdim).float() / (1 - p)
    yr = rq(h) @ rq(ps[2]).t() + ps[3] + (xr if add_residual else 0.0)
    yr.backward(dy.float())
    assert rel_err(y, yr) < tol and rel_err(x.grad, xr.grad) < tol
    for got, want in zip((w1, b1, w2, b2), ps):
        assert rel_err(got.grad, want.grad) < tol


def test_owned_gradients_match_autograd_accumulation():
    """With a FlatTrainer the backward kernels add straight into the flat gradient buffer (Linear weights / biases,
    LayerNorm gamma / beta); the result must equal what autograd's own accumulation produces, over two backward passes."""
    from acai_omr_b200 import layers, train

    def build():
        torch.manual_seed(11)
        m = torch.nn.Sequential()
        m.add_module("fc", layers.Linear(64, 128))
        m.add_module("ln", layers.LayerNorm(128))
        m.add_module("out", layers.Linear(128, 64))
        return m.to(DEV).train()

    def run(m, x):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            h = m.fc(x)
            h = m.ln(h, out_dtype=torch.bfloat16)
            y = m.out(h)
        (y.float() ** 2).mean().backward()

    x1, x2 = rnd(300, 64, seed=1), rnd(300, 64, seed=2)
    plain = build()
    run(plain, x1); run(plain, x2)                        # autograd accumulates .grad over the two passes
    owned = build()
    tr = train.FlatTrainer(owned, lr=1e-3)
    tr.zero_grad()
    run(owned, x1); run(owned, x2)
    for (n, p), (_, q) in zip(plain.named_parameters(), owned.named_parameters()):
        assert q.grad.data_ptr() >= tr.flat_grad.data_ptr() and rel_err(q.grad, p.grad) < 2e-3, n
