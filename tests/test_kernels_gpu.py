"""GPU: every C-ABI kernel against a plain-torch fp32 restatement of the same op, through ops.py
(= through the C ABI).  Shapes include ragged, empty, non-multiple-of-tile and odd head dims."""
import math

import pytest
import torch

from acai_omr_b200 import ops
from acai_omr_b200.layers import Packed
from oracle import vitomr_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


def close(a, b, rtol, atol):
    a, b = a.float(), b.float()
    err = (a - b).abs().max().item() if a.numel() else 0.0
    assert torch.allclose(a, b, rtol=rtol, atol=atol), f"max abs err {err}"


@pytest.mark.parametrize("M,N,K", [(1, 7, 5), (3, 227, 24), (8, 3072, 1024), (9, 64, 64), (65, 130, 33),
                                   (200, 96, 256), (0, 8, 8)])
@pytest.mark.parametrize("act", [ops.ACT_NONE, ops.ACT_GELU])
def test_linear_fp32(M, N, K, act):
    x, w, b, r = rnd(M, K, seed=1), rnd(N, K, seed=2, scale=0.2), rnd(N, seed=3), rnd(M, N, seed=4)
    y = ops.linear(x, w, b, act=act, residual=r)
    ref = x.double() @ w.double().t() + b.double()
    if act == ops.ACT_GELU:
        ref = torch.nn.functional.gelu(ref)
    ref = (ref + r.double()).float()
    close(y, ref, 1e-5, 2e-5)
    assert y.dtype == torch.float32 and y.shape == (M, N)


@pytest.mark.parametrize("M,N,K", [(5, 48, 40), (130, 200, 72), (4, 256, 128)])
def test_linear_bf16_operands(M, N, K):
    x, w, b = rnd(M, K, seed=1).bfloat16(), rnd(N, K, seed=2, scale=0.2).bfloat16(), rnd(N, seed=3)
    r = rnd(M, N, seed=4)
    y = ops.linear(x, w, b, act=ops.ACT_GELU, residual=r, out_dtype=torch.bfloat16)
    ref = torch.nn.functional.gelu(x.float() @ w.float().t() + b) + r
    close(y, ref, 1e-2, 1e-2)
    assert y.dtype == torch.bfloat16
    y32 = ops.linear(x, w, b, out_dtype=torch.float32)
    close(y32, x.float() @ w.float().t() + b, 1e-5, 1e-4)


def test_linear_strided_views():
    qkv = rnd(6, 3 * 16, seed=5)
    w = rnd(3 * 8, 16, seed=6)
    y = ops.linear(qkv[:, 16:32], w[8:16], None)
    close(y, qkv[:, 16:32] @ w[8:16].t(), 1e-5, 1e-5)


@pytest.mark.parametrize("rows,cols", [(1, 10), (7, 768), (33, 1024), (5, 48), (0, 16)])
@pytest.mark.parametrize("eps", [1e-5, 1e-6])
def test_layernorm(rows, cols, eps):
    x, r = rnd(rows, cols, seed=1, scale=3.0) + 0.5, rnd(rows, cols, seed=2)
    g, b = rnd(cols, seed=3) * 0.1 + 1.0, rnd(cols, seed=4) * 0.1
    y = ops.layernorm(x, g, b, eps, residual=r)
    ref = torch.nn.functional.layer_norm((x + r).double(), (cols,), g.double(), b.double(), eps).float()
    close(y, ref, 1e-5, 1e-5)
    y16, y32 = ops.layernorm(x.bfloat16(), g, b, eps, out_dtype=torch.bfloat16, out2_dtype=torch.float32)
    ref16 = torch.nn.functional.layer_norm(x.bfloat16().float(), (cols,), g, b, eps)
    close(y32, ref16, 1e-5, 1e-5)
    close(y16, ref16, 1e-2, 1e-2)


@pytest.mark.parametrize("rows,cols", [(1, 768), (9, 8), (5000, 768), (12345, 1024), (4737, 200), (300, 96), (4097, 768),
                                       (70001, 512), (4099, 64)])
@pytest.mark.parametrize("with_res", [False, True])
def test_layernorm_bf16_persistent(rows, cols, with_res):
    """bf16 -> bf16 LayerNorm: the persistent warp-per-row kernel with next-row prefetch (residual or < 4096 rows) and the
    per-warp TMA rings (>= 4096 rows, no residual; 1 / 2 / 4 rows per bulk copy, short last item, more items than warps)."""
    x = (rnd(rows, cols, seed=5, scale=2.0) + 0.3).bfloat16()
    r = rnd(rows, cols, seed=6).bfloat16() if with_res else None
    g, b = rnd(cols, seed=7) * 0.1 + 1.0, rnd(cols, seed=8) * 0.1
    y = ops.layernorm(x, g, b, 1e-5, residual=r, out_dtype=torch.bfloat16)
    xin = x.float() + (r.float() if with_res else 0.0)
    ref = torch.nn.functional.layer_norm(xin, (cols,), g, b, 1e-5)
    assert y.dtype == torch.bfloat16
    assert torch.equal(y, ref.bfloat16()) or float((y.float() - ref).abs().max()) <= 2.0 ** -7 * float(ref.abs().max())
    y32 = ops.layernorm(x, g, b, 1e-5, residual=r, out_dtype=torch.float32)          # generic kernel, same statistics
    close(y.float(), y32, 1.6e-2, 1.6e-2)


def ref_varlen(q, k, v, lens_q, lens_k, H, d, causal):
    out = torch.zeros(q.shape[0], H * d, device=q.device)
    oq = ok = 0
    for lq, lk in zip(lens_q, lens_k):
        if lq and lk:
            qq = q[oq:oq + lq].double().view(lq, H, d).transpose(0, 1)
            kk = k[ok:ok + lk].double().view(lk, H, d).transpose(0, 1)
            vv = v[ok:ok + lk].double().view(lk, H, d).transpose(0, 1)
            s = qq @ kk.transpose(1, 2) / math.sqrt(d)
            if causal:
                i = torch.arange(lq, device=q.device)[:, None] + (lk - lq)
                s = s.masked_fill(torch.arange(lk, device=q.device)[None, :] > i, float("-inf"))
            o = torch.softmax(s, -1) @ vv
            out[oq:oq + lq] = o.transpose(0, 1).reshape(lq, H * d).float()
        oq += lq
        ok += lk
    return out


@pytest.mark.parametrize("H,d", [(2, 6), (4, 8), (3, 32), (2, 64), (1, 100)])
@pytest.mark.parametrize("causal", [False, True])
def test_attn_varlen_self(H, d, causal):
    lens = [1, 70, 64, 5, 129]
    T, D = sum(lens), H * d
    qkv = rnd(T, 3 * D, seed=H * 10 + d)
    pk = Packed(lens, DEV)
    y = ops.attn_varlen(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], pk.cu, pk.cu, pk.max_len, H, d, causal=causal)
    ref = ref_varlen(qkv[:, :D].contiguous(), qkv[:, D:2 * D].contiguous(), qkv[:, 2 * D:].contiguous(), lens, lens, H, d, causal)
    close(y, ref, 1e-4, 2e-5)


def test_attn_varlen_cross_and_bf16():
    H, d = 4, 16
    lq, lk = [3, 40, 9], [17, 1, 66]
    q, kv = rnd(sum(lq), H * d, seed=1), rnd(sum(lk), 2 * H * d, seed=2)
    pq, pk = Packed(lq, DEV), Packed(lk, DEV)
    y = ops.attn_varlen(q, kv[:, :H * d], kv[:, H * d:], pq.cu, pk.cu, pq.max_len, H, d)
    ref = ref_varlen(q, kv[:, :H * d].contiguous(), kv[:, H * d:].contiguous(), lq, lk, H, d, False)
    close(y, ref, 1e-4, 2e-5)
    yb = ops.attn_varlen(q.bfloat16(), kv[:, :H * d].bfloat16(), kv[:, H * d:].bfloat16(), pq.cu, pk.cu, pq.max_len, H, d)
    refb = ref_varlen(q.bfloat16().float(), kv[:, :H * d].bfloat16().float(), kv[:, H * d:].bfloat16().float(), lq, lk, H, d, False)
    close(yb, refb, 2e-2, 2e-2)


@pytest.mark.parametrize("H,d,T", [(2, 6, 3), (16, 64, 300), (4, 32, 1), (3, 20, 77)])
def test_decode_attn_and_kv_append(H, d, T):
    B, Tmax = 3, T + 5
    kc = torch.zeros(B + 1, H, Tmax, d, device=DEV)
    vc = torch.zeros(B + 1, H, Tmax, d, device=DEV)
    ks, vs = rnd(B, H, T, d, seed=1), rnd(B, H, T, d, seed=2)
    for t in range(T):  # one position at a time, from a strided qkv-like buffer
        buf = torch.cat([ks[:, :, t].reshape(B, 1, H * d), vs[:, :, t].reshape(B, 1, H * d)], dim=-1)
        k_t, v_t = (c.view(B, H, 1, d) for c in buf.chunk(2, dim=-1))
        ops.kv_append(k_t, v_t, kc, vc, t)
    assert torch.equal(kc[:B, :, :T], ks) and torch.equal(vc[:B, :, :T], vs)
    assert kc[B].abs().sum() == 0 and kc[:, :, T:].abs().sum() == 0
    q = rnd(B, H * d, seed=3)
    mask = torch.zeros(B, T, dtype=torch.bool, device=DEV)
    if T > 2:
        mask[1, T // 2:] = True
    y = ops.decode_attn(q, kc[:B, :, :T], vc[:B, :, :T], T, H, d, key_padding_mask=mask)
    s = (q.view(B, H, 1, d).double() @ ks.double().transpose(-1, -2)) / math.sqrt(d)
    s = s.masked_fill(mask[:, None, None, :], float("-inf"))
    ref = (torch.softmax(s, -1) @ vs.double()).transpose(1, 2).reshape(B, H * d).float()
    close(y, ref, 1e-4, 2e-5)
    # multi-position append (KVCache.update with seq_len > 1)
    kc2, vc2 = torch.zeros_like(kc), torch.zeros_like(vc)
    ops.kv_append(ks, vs, kc2, vc2, 2)
    assert torch.equal(kc2[:B, :, 2:2 + T], ks)


def test_embed_argmax_pack_unpack():
    V, D, B, T = 227, 48, 3, 5
    table, pos = rnd(V, D, seed=1), rnd(9, D, seed=2)
    ids = torch.randint(0, V, (B, T), generator=torch.Generator().manual_seed(3)).to(DEV)
    y = ops.embed_tokens(ids, table, pos, torch.float32, period=T)
    assert torch.equal(y, table[ids] + pos[:T][None])
    y1 = ops.embed_tokens(ids[:, :1].int(), table, pos, torch.float32, pos_offset=4)
    assert torch.equal(y1, table[ids[:, :1]] + pos[4][None, None])

    logits = rnd(6, V, seed=4)
    logits[2, 100] = logits[2, 7] = logits[2].max() + 1.0  # tie -> lowest index
    nxt, lp = ops.argmax_logprob(logits)
    assert torch.equal(nxt, torch.argmax(logits, -1)) and nxt[2].item() == 7
    close(lp, torch.log_softmax(logits, -1).gather(-1, nxt[:, None]).squeeze(1), 1e-5, 1e-5)
    nb, lpb = ops.argmax_logprob(logits.bfloat16())
    assert torch.equal(nb, torch.argmax(logits.bfloat16(), -1))

    pk = Packed([2, 5, 1], DEV)
    rows = rnd(pk.total, D, seed=5)
    fill = rnd(D, seed=6)
    padded = ops.unpack_padded(rows, pk.cu, 3, 6, fill=fill)
    ref = O.pad_rows([rows[0:2], rows[2:7], rows[7:8]]).clone()
    ref = torch.cat([ref, torch.zeros(3, 1, D, device=DEV)], 1)
    m = Packed([2, 5, 1], DEV).mask(DEV, 6)
    ref[m] = fill
    assert torch.equal(padded, ref)
    assert torch.equal(ops.pack_padded(padded, pk.cu, pk.total), rows)
    assert ops.unpack_padded(rows, pk.cu, 3, 5)[0, 2:].abs().sum() == 0


@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_patch_embed_ragged(out_dtype):
    P, E = 4, 40
    shapes = [(8, 12), (16, 24), (4, 40), (24, 8), (9, 13)]  # last one has ragged borders that are dropped
    g = torch.Generator().manual_seed(0)
    imgs = [torch.rand(1, h, w, generator=g).to(DEV) for h, w in shapes]
    sd = {"projection.weight": rnd(E, P * P, seed=1, scale=0.3), "projection.bias": rnd(E, seed=2),
          "pos_embedding": rnd(6, 10, E, seed=3)}
    lens = [(h // P) * (w // P) for h, w in shapes]
    pk = Packed(lens, DEV)
    descs = ops.make_image_descs(imgs, P, DEV)
    y = ops.patch_embed(imgs, descs, pk.cu, pk.total, P, 1, sd["projection.weight"], sd["projection.bias"],
                        sd["pos_embedding"], False, out_dtype)
    cpu_sd = {k: v.cpu() for k, v in sd.items()}
    emb, mask = O.encoder_batchify([i.cpu() for i in imgs], cpu_sd, P, 6, 10)
    ref = emb[~mask]
    tol = (1e-5, 1e-5) if out_dtype == torch.float32 else (1e-2, 1e-2)
    close(y.cpu(), ref, *tol)
    # MAE-style gather of a subset of patches in shuffled order
    keep = [torch.tensor([3, 0, 5]), torch.tensor([20, 1]), torch.tensor([9]), torch.tensor([0, 11, 4, 7]), torch.tensor([2])]
    pk2 = Packed([len(k) for k in keep], DEV)
    idx = torch.cat(keep).int().to(DEV)
    y2 = ops.patch_embed(imgs, descs, pk2.cu, pk2.total, P, 1, sd["projection.weight"], sd["projection.bias"],
                         sd["pos_embedding"], False, torch.float32, patch_index=idx)
    ref2 = torch.cat([emb[i, k] for i, k in enumerate(keep)])
    close(y2.cpu(), ref2, 1e-5, 1e-5)


def test_patch_embed_tensor_core_path():
    """bf16 output + no interpolation -> im2col + tcgen05 GEMM with the position grid gathered in the epilogue."""
    P, E = 16, 192
    shapes = [(32, 64), (48, 160), (16, 16), (35, 50), (64, 33)]   # the last two have unaligned rows / dropped borders
    g = torch.Generator().manual_seed(5)
    imgs = [torch.rand(1, h, w, generator=g).to(DEV) for h, w in shapes]
    sd = {"projection.weight": rnd(E, P * P, seed=1, scale=0.06), "projection.bias": rnd(E, seed=2),
          "pos_embedding": rnd(5, 12, E, seed=3)}
    lens = [(h // P) * (w // P) for h, w in shapes]
    pk = Packed(lens, DEV)
    descs = ops.make_image_descs(imgs, P, DEV)
    n0 = ops.launch_count()
    y = ops.patch_embed(imgs, descs, pk.cu, pk.total, P, 1, sd["projection.weight"], sd["projection.bias"],
                        sd["pos_embedding"], False, torch.bfloat16)
    assert ops.launch_count() - n0 == 3
    cpu_sd = {k: v.cpu() for k, v in sd.items()}
    emb, mask = O.encoder_batchify([i.cpu() for i in imgs], cpu_sd, P, 5, 12)
    close(y.float().cpu(), emb[~mask], 2e-2, 2e-2)
    keep = [torch.tensor([7, 0, 5]), torch.tensor([20, 1, 29]), torch.tensor([0]), torch.tensor([0, 5, 4]), torch.tensor([2, 7])]
    pk2 = Packed([len(k) for k in keep], DEV)
    idx = torch.cat(keep).int().to(DEV)
    y2 = ops.patch_embed(imgs, descs, pk2.cu, pk2.total, P, 1, sd["projection.weight"], sd["projection.bias"],
                         sd["pos_embedding"], False, torch.bfloat16, patch_index=idx)
    ref2 = torch.cat([emb[i, k] for i, k in enumerate(keep)])
    close(y2.float().cpu(), ref2, 2e-2, 2e-2)


def test_patch_embed_interpolated_pe():
    P, E = 2, 8
    pe = rnd(4, 6, E, seed=1)
    imgs = [torch.rand(1, 12, 6).to(DEV), torch.rand(1, 4, 20).to(DEV), torch.rand(1, 4, 4).to(DEV)]
    w, b = rnd(E, 4, seed=2), rnd(E, seed=3)
    lens = [18, 20, 4]
    pk = Packed(lens, DEV)
    y = ops.patch_embed(imgs, ops.make_image_descs(imgs, P, DEV), pk.cu, pk.total, P, 1, w, b, pe, True, torch.float32)
    sd = {"projection.weight": w.cpu(), "projection.bias": b.cpu(), "pos_embedding": pe.cpu()}
    emb, mask = O.encoder_batchify([i.cpu() for i in imgs], sd, P, 4, 6, interpolate=True)
    close(y.cpu(), emb[~mask], 1e-5, 1e-5)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("lens", [[1024, 968, 972], [1, 127, 128, 129, 300], [64], [257, 5]])
def test_attn_varlen_tcgen05_d64(lens, causal):
    """bf16, head_dim 64 -> the tcgen05 kernel (attn_tc.cu); reference = fp32 math on the same bf16 inputs."""
    H, d = 3, 64
    T, D = sum(lens), H * d
    qkv = rnd(T, 3 * D, seed=len(lens) + 7).bfloat16()
    pk = Packed(lens, DEV)
    y = ops.attn_varlen(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], pk.cu, pk.cu, pk.max_len, H, d, causal=causal)
    ref = ref_varlen(qkv[:, :D].float().contiguous(), qkv[:, D:2 * D].float().contiguous(),
                     qkv[:, 2 * D:].float().contiguous(), lens, lens, H, d, causal)
    assert y.dtype == torch.bfloat16
    close(y, ref, 2e-2, 2e-2)
    y32 = ops.attn_varlen(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], pk.cu, pk.cu, pk.max_len, H, d, causal=causal,
                          out_dtype=torch.float32)
    close(y32, ref, 1e-2, 1e-2)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("kind", ["ramp", "jump", "late_spike"])
def test_attn_varlen_tcgen05_growing_row_maximum(kind, causal, monkeypatch):
    """Scores whose running row maximum keeps moving: every 32-key block beats the previous maximum by more than 2^8 ("ramp": the
    O / l rescale runs in every block), one block jumps by ~2^70 ("jump": the speculative path must recompute that block under its own
    maximum instead of overflowing), the last keys dominate ("late_spike").  Both softmax arrangements (reference maximum of the
    block itself / of the previous block, ACAI_ATTN_SPEC) against fp32 math, and the log-sum-exp the backward reads."""
    H, d = 2, 64
    lens = [300, 129, 64]
    T, D = sum(lens), H * d
    q = torch.ones(T, D, device=DEV)
    k = torch.zeros(T, D, device=DEV)
    v = rnd(T, D, seed=31)
    off = 0
    for L in lens:
        pos = torch.arange(L, device=DEV, dtype=torch.float32)
        if kind == "ramp":
            b = 0.05 * pos                                      # score = 64 b / 8 = 0.4 nats per key: 12.8 nats per 32-key block
        elif kind == "jump":
            b = torch.where(pos >= min(200, L // 2), 6.25, 0.01 * pos)   # one step of ~50 nats
        else:
            b = torch.where(pos >= L - 3, 3.0 + 0.1 * (pos - L), 0.0)
        k[off:off + L] = b[:, None]
        off += L
    q, k, v = q.bfloat16(), k.bfloat16(), v.bfloat16()
    pk = Packed(lens, DEV)
    ref = ref_varlen(q.float(), k.float(), v.float(), lens, lens, H, d, causal)
    outs = []
    for spec in ("0", "1"):
        monkeypatch.setenv("ACAI_ATTN_SPEC", spec)              # (read per call by the library)
        y, lse = ops.attn_varlen(q, k, v, pk.cu, pk.cu, pk.max_len, H, d, causal=causal, out_dtype=torch.float32, return_lse=True)
        assert bool(torch.isfinite(y).all()) and bool(torch.isfinite(lse).all())
        close(y, ref, 2e-2, 2e-2)
        outs.append(y)


def test_attn_varlen_tcgen05_cross():
    H, d = 2, 64
    lq, lk = [200, 3, 130], [1000, 140, 17]
    q, kv = rnd(sum(lq), H * d, seed=1).bfloat16(), rnd(sum(lk), 2 * H * d, seed=2).bfloat16()
    pq, pk = Packed(lq, DEV), Packed(lk, DEV)
    y = ops.attn_varlen(q, kv[:, :H * d], kv[:, H * d:], pq.cu, pk.cu, pq.max_len, H, d)
    ref = ref_varlen(q.float(), kv[:, :H * d].float().contiguous(), kv[:, H * d:].float().contiguous(), lq, lk, H, d, False)
    close(y, ref, 2e-2, 2e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_decode_attn_paged_self_and_cross(dtype):
    """acai_decode_attn_paged: paged self cache with append driven by a device step counter, and the packed memory
    cache with per-sequence lengths; vs fp32 math on the same (rounded) values."""
    B, H, d, max_len = 3, 4, 64, 700
    D = H * d
    pps = (max_len + 63) // 64
    g = torch.Generator().manual_seed(0)
    perm = torch.randperm(B * pps, generator=g).int().view(B, pps).to(DEV)      # scattered pages
    pool = torch.zeros(B * pps * 64, 2 * D, dtype=dtype, device=DEV)
    ws = ops.decode_workspace(B, H, max_len, DEV)
    step = torch.ones(1, dtype=torch.int32, device=DEV)
    hist = []
    tol = (1e-4, 2e-5) if dtype == torch.float32 else (2e-2, 2e-2)
    for t in list(range(1, 6)) + [255, 256, 257, 600]:
        # fill history up to t-1 cached tokens directly, then run one step that appends token t-1
        while len(hist) < t - 1:
            row = rnd(B, 2 * D, seed=100 + len(hist)).to(dtype)
            pos = len(hist)
            rows = perm[:, pos // 64].long() * 64 + pos % 64
            pool[rows] = row
            hist.append(row)
        step.fill_(t)
        qkv = rnd(B, 3 * D, seed=500 + t).to(dtype)
        out = ops.decode_attn_paged(qkv[:, :D], qkv[:, D:], pool, H, d, max_len, ws, torch.float32,
                                    page_table=perm, step=step, len_offset=-1)
        hist.append(qkv[:, D:].clone())
        kv = torch.stack(hist, dim=1).float()                                        # (B, t, 2D)
        k, v = kv[..., :D].view(B, t, H, d).transpose(1, 2), kv[..., D:].view(B, t, H, d).transpose(1, 2)
        s = (qkv[:, :D].float().view(B, H, 1, d).double() @ k.double().transpose(-1, -2)) / math.sqrt(d)
        ref = (torch.softmax(s, -1) @ v.double()).transpose(1, 2).reshape(B, D).float()
        close(out, ref, *tol)
        rows = perm[:, (t - 1) // 64].long() * 64 + (t - 1) % 64
        assert torch.equal(pool[rows], qkv[:, D:])                                   # the append landed in its page
    # packed memory cache, ragged lengths, no append
    lens = [300, 1, 257]
    mp = Packed(lens, DEV)
    mem = rnd(sum(lens), 2 * D, seed=9).to(dtype)
    q = rnd(B, D, seed=10).to(dtype)
    wsc = ops.decode_workspace(B, H, max(lens), DEV)
    out = ops.decode_attn_paged(q, None, mem, H, d, max(lens), wsc, torch.float32, seq_start=mp.cu[:-1].contiguous(),
                                kv_lens=torch.tensor(lens, dtype=torch.int32, device=DEV))
    o = 0
    for b, L in enumerate(lens):
        kv = mem[o:o + L].float()
        k, v = kv[:, :D].view(L, H, d).transpose(0, 1), kv[:, D:].view(L, H, d).transpose(0, 1)
        s = (q[b].float().view(H, 1, d).double() @ k.double().transpose(-1, -2)) / math.sqrt(d)
        ref = (torch.softmax(s, -1) @ v.double()).reshape(D).float()
        close(out[b], ref, *tol)
        o += L


@pytest.mark.parametrize("M", [1, 2, 5, 8, 13, 16, 32, 33])
@pytest.mark.parametrize("N,K", [(1024, 1024), (227, 1024), (4096, 1024), (1024, 4096), (40, 72)])
def test_linear_bf16_decode_step_shapes(M, N, K):
    """M <= 32 takes the weight-streaming GEMV (one warp per output column), M = 33 the tensor-core tile: same results."""
    x, w = rnd(M, K, seed=71).bfloat16(), rnd(N, K, seed=72, scale=0.05).bfloat16()
    b, r = rnd(N, seed=73), rnd(M, N, seed=74).bfloat16()
    for act in (ops.ACT_NONE, ops.ACT_GELU):
        res = None if act else r
        y = ops.linear(x, w, b, act=act, residual=res, out_dtype=torch.bfloat16)
        ref = x.double() @ w.double().t() + b.double()
        if act:
            ref = torch.nn.functional.gelu(ref)
        else:
            ref = ref + r.double()
        close(y, ref.float(), 1e-2, 2e-2)
    # strided activation rows (a column slice of a fused QKV buffer)
    big = rnd(M, 3 * K, seed=75).bfloat16()
    y = ops.linear(big[:, K:2 * K], w, b, out_dtype=torch.float32)
    close(y, (big[:, K:2 * K].double() @ w.double().t() + b.double()).float(), 1e-3, 2e-3)


def test_nvtx_ranges_around_entry_points(monkeypatch):
    """ACAI_NVTX=1: every C-ABI call is wrapped in an NVTX range named after its entry point (push / pop balanced, results unchanged)."""
    pushed, popped = [], []
    monkeypatch.setattr(ops, "NVTX", True)
    monkeypatch.setattr(torch.cuda.nvtx, "range_push", lambda name: pushed.append(name))
    monkeypatch.setattr(torch.cuda.nvtx, "range_pop", lambda: popped.append(1))
    x = torch.randn(33, 64, device="cuda")
    g, b = torch.ones(64, device="cuda"), torch.zeros(64, device="cuda")
    y = ops.layernorm(x, g, b, 1e-5)
    assert pushed and len(pushed) == len(popped) and pushed[0].startswith("acai_layernorm")
    assert torch.allclose(y, torch.nn.functional.layer_norm(x, (64,), g, b, 1e-5), atol=1e-5)
