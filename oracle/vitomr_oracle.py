"""CPU oracle for the ViTOMR encoder / LMX decoder hot path.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
legs may import it.  The product path (`acai_omr_b200`) never does.

It is a plain-`torch`, fp32, explicit-math restatement (matmul / softmax /
mean / var spelled out; no `nn.Transformer*`, no SDPA, no nested tensors) of the
reference algorithm in `acai_omr/models/models.py` and
`acai_omr/models/kv_caching.py` (jsnchon/acai-omr @ 5819ae3).  Every function
cites the reference `file:line` it follows.  It works directly on a reference
`state_dict` (same key names), so weights can be moved between the reference,
the oracle and the CUDA modules without renaming.

Parity pin: `tests/golden/golden_small.pt` holds outputs of the UNMODIFIED
reference modules (imported from /root/reference in the build container by
`tests/golden/make_golden.py`, torch 2.11 CPU fp32).  `tests/test_oracle.py`
checks this oracle against every entry, plus the reference's own known-answer
tests (`tests/test_mae.py:169-180` MAELoss == 10.583329200744629, the exact
`batchify` values of `tests/test_vitomr.py:27-49`, `tests/test_mae.py:8-24,57-79,
89-122`, `tests/test_vitomr.py:151-172`).  Full-model logits / token sequences
have no stored fixture in the reference itself ("pinned by live oracle").
"""
from __future__ import annotations

import contextlib
import math
from dataclasses import dataclass, field

import torch

PAD_TOKEN, BOS_TOKEN, EOS_TOKEN = "<pad>", "<bos>", "<eos>"  # acai_omr/config.py:12-14


# --------------------------------------------------------------------------- #
# precision mode
# --------------------------------------------------------------------------- #
# The reference's production inference runs the transition head and the cached greedy loop under
# `torch.autocast(dtype=torch.bfloat16)` with bf16 KV caches, the encoder in fp32 outside it
# (acai_omr/inference/vitomr_inference.py:79-86,94).  `autocast_bf16()` restates what autocast does to the ops on that
# path: `F.linear` / `nn.Linear` and `F.scaled_dot_product_attention` take bf16 operands (fp32 accumulation, ONE rounding
# of the result to bf16); GELU runs in its input's dtype (bf16 after a Linear); residual adds promote to the fp32 stream
# (the token embedding + position rows are fp32 parameters), so every LayerNorm sees and returns fp32; the logits come
# out of `unembed` in bf16 and argmax / log_softmax read those.  Pinned by tests/golden/golden_bf16.pt (the unmodified
# reference under `torch.autocast("cpu", torch.bfloat16)`; tests/golden/make_golden_bf16.py).
_AUTOCAST_BF16 = False
_ROUNDED = {}          # bf16-rounded copies of weight tensors, alive for the duration of one autocast_bf16() block


@contextlib.contextmanager
def autocast_bf16(enabled=True):
    global _AUTOCAST_BF16
    prev, _AUTOCAST_BF16 = _AUTOCAST_BF16, bool(enabled)
    try:
        yield
    finally:
        _AUTOCAST_BF16 = prev
        if not prev:
            _ROUNDED.clear()


def _r(x):
    """Round to bf16, keep computing in fp32."""
    return x.to(torch.bfloat16).to(torch.float32)


def _rw(w):
    """`_r` of a weight (slice), cached by storage address + shape: a decode loop reads the same weights every step."""
    key = (w.data_ptr(), tuple(w.shape), tuple(w.stride()))
    hit = _ROUNDED.get(key)
    if hit is None:
        hit = _ROUNDED[key] = (_r(w), w)      # the source tensor is kept alive so that its address cannot be reused
    return hit[0]


# --------------------------------------------------------------------------- #
# primitives
# --------------------------------------------------------------------------- #
def linear(x, w, b=None):
    """y = x W^T + b  (torch `F.linear`; used at kv_caching.py:193,215,244)."""
    if _AUTOCAST_BF16:
        y = _r(x) @ _rw(w).t()
        if b is not None:
            y = y + _rw(b)
        return y.to(torch.bfloat16)
    y = x @ w.t()
    return y if b is None else y + b


def gelu(x):
    """Exact-erf GELU (`activation="gelu"`, models.py:31; `nn.GELU()`, models.py:657)."""
    xf = x.float()
    return (0.5 * xf * (1.0 + torch.erf(xf * (1.0 / math.sqrt(2.0))))).to(x.dtype)


def layer_norm(x, w, b, eps):
    """LayerNorm over the last dim with biased variance (torch semantics; fp32 statistics for any input dtype)."""
    xf = x.float()
    mu = xf.mean(dim=-1, keepdim=True)
    var = ((xf - mu) ** 2).mean(dim=-1, keepdim=True)
    return ((xf - mu) / torch.sqrt(var + eps) * w + b).to(x.dtype)


def patchify(img, P):
    """`nn.Unfold(P, P)` then transpose (models.py:48,52): (C,H,W) -> (L, C*P*P).

    Patches are row-major over (h_p, w_p); inside a patch the feature order is
    (channel, row-in-patch, col-in-patch).  Pixels beyond h_p*P / w_p*P are dropped.
    """
    C, H, W = img.shape
    hp, wp = H // P, W // P
    x = img[:, : hp * P, : wp * P].reshape(C, hp, P, wp, P)
    return x.permute(1, 3, 0, 2, 4).reshape(hp * wp, C * P * P)


def interpolate_pe(pe, hp, wp):
    """Bilinear, align_corners=False resize of the (Hm,Wm,E) PE grid (models.py:291-302).

    Restated from the `F.interpolate` definition: src = (dst + 0.5) * scale - 0.5,
    clamped at 0; neighbours clamped at the border.
    """
    Hm, Wm, _ = pe.shape

    def axis(n_out, n_in):
        scale = n_in / n_out
        src = (torch.arange(n_out, dtype=torch.float32) + 0.5) * scale - 0.5
        src = src.clamp(min=0.0)
        i0 = src.floor().long().clamp(max=n_in - 1)
        i1 = (i0 + 1).clamp(max=n_in - 1)
        lam = (src - i0.float()).clamp(0.0, 1.0)
        return i0, i1, lam

    r0, r1, lr = axis(hp, Hm)
    c0, c1, lc = axis(wp, Wm)
    pe = pe.float().cpu()
    top = pe[r0][:, c0] * (1 - lc)[None, :, None] + pe[r0][:, c1] * lc[None, :, None]
    bot = pe[r1][:, c0] * (1 - lc)[None, :, None] + pe[r1][:, c1] * lc[None, :, None]
    return top * (1 - lr)[:, None, None] + bot * lr[:, None, None]


def mha(x_q, x_kv, in_w, in_b, out_w, out_b, H, key_pad_mask=None, causal=False):
    """`nn.MultiheadAttention` forward, batch_first (kv_caching.py:193-202,212-216,241-250 spell out
    the `[Wq;Wk;Wv]` row layout and the contiguous head split).  Masks: True = ignore."""
    B, Tq, D = x_q.shape
    Tk = x_kv.shape[1]
    d = D // H
    q = linear(x_q, in_w[:D], in_b[:D]).view(B, Tq, H, d).transpose(1, 2)
    k = linear(x_kv, in_w[D : 2 * D], in_b[D : 2 * D]).view(B, Tk, H, d).transpose(1, 2)
    v = linear(x_kv, in_w[2 * D :], in_b[2 * D :]).view(B, Tk, H, d).transpose(1, 2)
    s = (q.float() @ k.float().transpose(-1, -2)) / math.sqrt(d)
    if causal:
        tri = torch.triu(torch.ones(Tq, Tk, dtype=torch.bool, device=s.device), diagonal=1)
        s = s.masked_fill(tri, float("-inf"))
    if key_pad_mask is not None:
        s = s.masked_fill(key_pad_mask[:, None, None, :], float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = (p @ v.float()).to(q.dtype).transpose(1, 2).reshape(B, Tq, D)
    return linear(o, out_w, out_b)


# --------------------------------------------------------------------------- #
# encoder family  (models.py:14-96, 290-376)
# --------------------------------------------------------------------------- #
def encoder_layer(x, sd, pre, H, key_pad_mask):
    """Post-norm `nn.TransformerEncoderLayer` (models.py:31): x=LN1(x+SA(x)); x=LN2(x+FF(x))."""
    a = mha(x, x, sd[pre + "self_attn.in_proj_weight"], sd[pre + "self_attn.in_proj_bias"],
            sd[pre + "self_attn.out_proj.weight"], sd[pre + "self_attn.out_proj.bias"], H, key_pad_mask)
    x = layer_norm(x + a, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], 1e-5)
    f = linear(gelu(linear(x, sd[pre + "linear1.weight"], sd[pre + "linear1.bias"])),
               sd[pre + "linear2.weight"], sd[pre + "linear2.bias"])
    return layer_norm(x + f, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], 1e-5)


def count_layers(sd, stack):
    n = 0
    while f"{stack}.layers.{n}.norm1.weight" in sd:
        n += 1
    return n


def encoder_stack(x, sd, stack, H, key_pad_mask):
    """`nn.TransformerEncoder` (models.py:30-34): N layers then optional final LN eps 1e-6."""
    for i in range(count_layers(sd, stack)):
        x = encoder_layer(x, sd, f"{stack}.layers.{i}.", H, key_pad_mask)
    if f"{stack}.norm.weight" in sd:
        x = layer_norm(x, sd[f"{stack}.norm.weight"], sd[f"{stack}.norm.bias"], 1e-6)
    return x


def pad_rows(seqs, pad_value=0.0):
    """nested(jagged) -> to_padded_tensor (models.py:55-56)."""
    Lm = max(s.shape[0] for s in seqs)
    out = seqs[0].new_full((len(seqs), Lm) + tuple(seqs[0].shape[1:]), pad_value)
    for i, s in enumerate(seqs):
        out[i, : s.shape[0]] = s
    return out


def attention_mask(seq_lens, max_len):
    """create_attention_mask (models.py:70-73): True = padding."""
    return torch.arange(max_len)[None, :] >= torch.tensor(seq_lens)[:, None]


def encoder_batchify(imgs, sd, P, pe_max_h, pe_max_w, interpolate=False, identity_projection=False):
    """Encoder.batchify (models.py:36-66) / OMREncoder.batchify (models.py:304-332)."""
    pe = sd["pos_embedding"]
    E = pe.shape[-1]
    embs, lens = [], []
    for t in imgs:
        hp, wp = t.shape[-2] // P, t.shape[-1] // P
        if hp > pe_max_h or wp > pe_max_w:
            if not interpolate:
                raise ValueError(f"{hp} x {wp} image is too large for max positional embedding grid "
                                 f"of shape {pe_max_h} x {pe_max_w}")
            pe_slice = interpolate_pe(pe, hp, wp).reshape(-1, E)
        else:
            pe_slice = pe[:hp, :wp, :].reshape(-1, E)
        patches = patchify(t, P)
        lens.append(patches.shape[0])
        # the projection is applied to the PADDED batch in the reference, so pad rows get the bias;
        # they are never read as valid data.  We keep that: pad rows = bias (+0 PE).
        embs.append((patches, pe_slice))
    padded = pad_rows([p for p, _ in embs])
    x = padded if identity_projection else linear(padded, sd["projection.weight"], sd["projection.bias"])
    x = x + pad_rows([pe_ for _, pe_ in embs])
    return x, attention_mask(lens, x.shape[1])


def encoder_forward(imgs, sd, cfg, fast_path_pad_fill=True):
    """Encoder.forward (models.py:76-79) / FineTuneOMREncoder.forward (models.py:362-367).

    `cfg`: dict(patch_size, pe_max_height, pe_max_width, num_heads, interpolate, stacks=[...]).
    With `fast_path_pad_fill` the padded rows of the result hold the final norm's bias, which is
    what torch's eval-mode nested-tensor fast path returns: it un-pads, runs the
    layers, re-pads with 0.0 and only then applies the final LayerNorm.
    """
    x, mask = encoder_batchify(imgs, sd, cfg["patch_size"], cfg["pe_max_height"], cfg["pe_max_width"],
                               interpolate=cfg.get("interpolate", False))
    stacks = cfg["stacks"]
    for stack in stacks:
        x = encoder_stack(x, sd, stack, cfg["num_heads"], mask)
    if fast_path_pad_fill and f"{stacks[-1]}.norm.bias" in sd:
        x = torch.where(mask[..., None], sd[f"{stacks[-1]}.norm.bias"].expand_as(x), x)
    return x, mask


def transition_head(x, sd, pre="transition_head."):
    """Linear -> GELU -> (Dropout, eval: identity) -> Linear (models.py:655-660)."""
    return linear(gelu(linear(x, sd[pre + "0.weight"], sd[pre + "0.bias"])), sd[pre + "3.weight"], sd[pre + "3.bias"])


# --------------------------------------------------------------------------- #
# LMX decoder  (models.py:378-528, kv_caching.py)
# --------------------------------------------------------------------------- #
def decoder_layer(x, memory, sd, pre, H, causal, tgt_pad_mask, mem_pad_mask):
    """Post-norm `nn.TransformerDecoderLayer` (models.py:423; cached twin kv_caching.py:190-223)."""
    a = mha(x, x, sd[pre + "self_attn.in_proj_weight"], sd[pre + "self_attn.in_proj_bias"],
            sd[pre + "self_attn.out_proj.weight"], sd[pre + "self_attn.out_proj.bias"], H, tgt_pad_mask, causal)
    x = layer_norm(x + a, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], 1e-5)
    c = mha(x, memory, sd[pre + "multihead_attn.in_proj_weight"], sd[pre + "multihead_attn.in_proj_bias"],
            sd[pre + "multihead_attn.out_proj.weight"], sd[pre + "multihead_attn.out_proj.bias"], H, mem_pad_mask)
    x = layer_norm(x + c, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], 1e-5)
    f = linear(gelu(linear(x, sd[pre + "linear1.weight"], sd[pre + "linear1.bias"])),
               sd[pre + "linear2.weight"], sd[pre + "linear2.bias"])
    return layer_norm(x + f, sd[pre + "norm3.weight"], sd[pre + "norm3.bias"], 1e-5)


def decoder_stack(x, memory, sd, H, causal=True, tgt_pad_mask=None, mem_pad_mask=None, stack="decoder_blocks"):
    for i in range(count_layers(sd, stack)):
        x = decoder_layer(x, memory, sd, f"{stack}.layers.{i}.", H, causal, tgt_pad_mask, mem_pad_mask)
    if f"{stack}.norm.weight" in sd:
        x = layer_norm(x, sd[f"{stack}.norm.weight"], sd[f"{stack}.norm.bias"], 1e-6)
    return x


def decoder_forward(tokens, memory, sd, H, lmx_pad_mask=None, mem_pad_mask=None):
    """OMRDecoder.forward / .generate (models.py:445-500): embed + PE[:T], causal stack, unembed."""
    T = tokens.shape[1]
    x = sd["vocab_embedding.weight"][tokens.long()] + sd["pos_embedding"][:T][None]
    h = decoder_stack(x, memory, sd, H, True, lmx_pad_mask, mem_pad_mask)
    return linear(h, sd["unembed.weight"], sd["unembed.bias"])


@dataclass
class CachedDecoderState:
    """KVCache + MemoryCache contents (kv_caching.py:5-109, 227-256), one entry per layer."""
    k_self: list = field(default_factory=list)
    v_self: list = field(default_factory=list)
    k_cross: list = field(default_factory=list)
    v_cross: list = field(default_factory=list)


def prepare_caches(memory, sd, H, stack="decoder_blocks"):
    """CachedTransformerDecoder.prepare_caches (kv_caching.py:283-287) + MemoryCache (:235-253)."""
    B, S, D = memory.shape
    d = D // H
    st = CachedDecoderState()
    for i in range(count_layers(sd, stack)):
        pre = f"{stack}.layers.{i}.multihead_attn."
        kv = linear(memory, sd[pre + "in_proj_weight"][D:], sd[pre + "in_proj_bias"][D:])
        k, v = kv.chunk(2, dim=-1)
        # .float().contiguous(): in bf16 mode the caches hold bf16-rounded VALUES in fp32 containers (exact), laid out
        # head-major, so that the per-step attention neither re-converts nor re-strides the whole cache
        st.k_cross.append(k.view(B, S, H, d).transpose(1, 2).float().contiguous())
        st.v_cross.append(v.view(B, S, H, d).transpose(1, 2).float().contiguous())
        st.k_self.append(torch.zeros(B, H, 0, d))
        st.v_self.append(torch.zeros(B, H, 0, d))
    return st


def _attend_one(q, K, V, mask=None):
    """CachedMultiheadAttention.cached_forward's SDPA (kv_caching.py:123-136); q:(B,H,1,d)."""
    s = (q.float() @ K.float().transpose(-1, -2)) / math.sqrt(q.shape[-1])
    if mask is not None:
        s = s.masked_fill(mask[:, None, None, :], float("-inf"))
    o = (torch.softmax(s, dim=-1) @ V.float()).to(q.dtype)
    return o.transpose(1, 2).flatten(start_dim=-2)


def cached_decoder_step(x, st, sd, H, mem_pad_mask=None, stack="decoder_blocks"):
    """CachedTransformerDecoder.cached_generate (kv_caching.py:292-302); x:(B,1,D)."""
    B, _, D = x.shape
    d = D // H
    for i in range(count_layers(sd, stack)):
        pre = f"{stack}.layers.{i}."
        qkv = linear(x, sd[pre + "self_attn.in_proj_weight"], sd[pre + "self_attn.in_proj_bias"])
        q, k, v = (t.view(B, H, 1, d) for t in qkv.chunk(3, dim=-1))  # kv_caching.py:194-202 (a view, not a transpose)
        st.k_self[i] = torch.cat([st.k_self[i], k.float()], dim=2)
        st.v_self[i] = torch.cat([st.v_self[i], v.float()], dim=2)
        a = linear(_attend_one(q, st.k_self[i], st.v_self[i]),
                   sd[pre + "self_attn.out_proj.weight"], sd[pre + "self_attn.out_proj.bias"])
        x = layer_norm(x + a, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], 1e-5)
        qc = linear(x, sd[pre + "multihead_attn.in_proj_weight"][:D], sd[pre + "multihead_attn.in_proj_bias"][:D])
        c = linear(_attend_one(qc.view(B, H, 1, d), st.k_cross[i], st.v_cross[i], mem_pad_mask),
                   sd[pre + "multihead_attn.out_proj.weight"], sd[pre + "multihead_attn.out_proj.bias"])
        x = layer_norm(x + c, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], 1e-5)
        f = linear(gelu(linear(x, sd[pre + "linear1.weight"], sd[pre + "linear1.bias"])),
                   sd[pre + "linear2.weight"], sd[pre + "linear2.bias"])
        x = layer_norm(x + f, sd[pre + "norm3.weight"], sd[pre + "norm3.bias"], 1e-5)
    if f"{stack}.norm.weight" in sd:
        x = layer_norm(x, sd[f"{stack}.norm.weight"], sd[f"{stack}.norm.bias"], 1e-6)
    return x


def cached_generate(token_t, time_step, st, sd, H, mem_pad_mask=None):
    """OMRDecoder.cached_generate (models.py:518-528)."""
    if time_step >= sd["pos_embedding"].shape[0]:
        raise RuntimeError(f"{time_step + 1} decoding steps is too long for max sequence length "
                           f"of {sd['pos_embedding'].shape[0]}")
    x = sd["vocab_embedding.weight"][token_t.long()] + sd["pos_embedding"][time_step][None, None]
    h = cached_decoder_step(x, st, sd, H, mem_pad_mask)
    return linear(h, sd["unembed.weight"], sd["unembed.bias"])


def create_inference_mask(seqs, eos_idx):
    """ViTOMR.create_inference_mask (models.py:550-559)."""
    eos = seqs == eos_idx
    seen = eos.int().cumsum(dim=-1)
    return (seen == 0) | (eos & (seen == 1))


def cached_greedy_generate(memory, sd, H, mem_pad_mask=None, max_len=1536, pad_idx=1, bos_idx=0, eos_idx=2):
    """ViTOMR.cached_greedy_generate (models.py:600-615) incl. the time_step = t quirk (:576)."""
    B = memory.shape[0]
    st = prepare_caches(memory, sd, H)
    seqs = torch.full((B, max_len), pad_idx, dtype=torch.long)
    seqs[:, 0] = bos_idx
    logp = torch.zeros(B, max_len, dtype=torch.float32)
    finished = torch.zeros(B, dtype=torch.bool)
    all_logits = []
    for t in range(1, max_len):
        logits = cached_generate(seqs[:, t - 1 : t], t, st, sd, H, mem_pad_mask).squeeze(1)
        all_logits.append(logits)
        nxt = torch.argmax(logits, dim=-1)
        lp = torch.log_softmax(logits.float(), dim=-1).gather(-1, nxt[:, None]).squeeze(1)
        seqs[:, t] = nxt
        logp[:, t] = lp
        finished |= nxt == eos_idx
        if bool(finished.all()):
            break
    m = create_inference_mask(seqs, eos_idx)  # mask_and_clip_seqs, models.py:585-596
    seqs = seqs.masked_fill(~m, pad_idx)
    logp = logp.masked_fill(~m, 0.0)
    n = int(m.sum(dim=-1).max())
    return seqs[:, :n], logp[:, :n], m[:, :n], all_logits


def cached_forced_logits(memory, sd, H, tokens, mem_pad_mask=None):
    """The cached decoder step of cached_greedy_generate driven by GIVEN tokens: logits[:, t-1] is what the loop sees at
    step t (input tokens[:, t-1], `time_step = t`, models.py:575-577) -> (B, T-1, V).  Lets a second implementation be
    compared step by step on the same prefix, where a free-running comparison stops meaning anything after the first
    near-tie flips."""
    st = prepare_caches(memory, sd, H)
    out = []
    for t in range(1, tokens.shape[1]):
        out.append(cached_generate(tokens[:, t - 1 : t], t, st, sd, H, mem_pad_mask).squeeze(1))
    return torch.stack(out, dim=1)


def batchify_and_split_lmx_seqs(lmx_seqs, pad_idx):
    """models.py:531-540."""
    padded = pad_rows([s.long() for s in lmx_seqs], pad_idx)
    inp, tgt = padded[:, :-1], padded[:, 1:]
    return inp, tgt, inp == pad_idx


def ce_loss(pred, target, pad_idx):
    """OMRCELoss (models.py:784-796): mean NLL over non-pad targets."""
    lp = torch.log_softmax(pred.reshape(-1, pred.shape[-1]).float(), dim=-1)
    t = target.reshape(-1).long()
    keep = t != pad_idx
    nll = -lp.gather(-1, t[:, None]).squeeze(1)
    return (nll * keep).sum() / keep.sum()


def teacher_forced_forward(batch, sd, cfg):
    """TeacherForcedViTOMR.forward (models.py:722-736), eval/dropout-0 semantics, train-mode encoder
    (no pad-row fill; pad rows are masked everywhere downstream)."""
    imgs, seqs = zip(*batch)
    enc_sd = {k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")}
    dec_sd = {k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}
    lat, lat_mask = encoder_forward(list(imgs), enc_sd, cfg["encoder"], fast_path_pad_fill=False)
    lat = transition_head(lat, sd)
    inp, tgt, lmx_mask = batchify_and_split_lmx_seqs(list(seqs), cfg["pad_idx"])
    pred = decoder_forward(inp, lat, dec_sd, cfg["decoder_heads"], lmx_mask, lat_mask)
    return pred, tgt


# --------------------------------------------------------------------------- #
# MAE  (models.py:100-288)
# --------------------------------------------------------------------------- #
def mask_sequence(noise, mask_ratio):
    """MAEEncoder.mask_sequence index bookkeeping (models.py:106-125) for a given noise vector."""
    L = noise.shape[0]
    len_keep = int(L * (1 - mask_ratio))
    ids_shuffle = torch.argsort(noise)
    ids_restore = torch.argsort(ids_shuffle)
    ids_keep = ids_shuffle[:len_keep]
    seq_mask = torch.ones(L, dtype=torch.int)
    seq_mask[:len_keep] = 0
    return ids_keep, ids_restore, seq_mask.index_select(0, ids_restore), len_keep


def mae_forward(batch, sd, cfg, noises):
    """MAE.forward (models.py:249-269) with the per-image `torch.rand` noise supplied by the caller."""
    xs, ys = zip(*batch)
    P = cfg["patch_size"]
    enc = {k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")}
    E = enc["pos_embedding"].shape[-1]
    kept, kept_pe, keep_lens, full_lens, seq_masks, restores, dims = [], [], [], [], [], [], []
    for t, noise in zip(xs, noises):
        hp, wp = t.shape[-2] // P, t.shape[-1] // P
        if hp > cfg["pe_max_height"] or wp > cfg["pe_max_width"]:
            raise ValueError("image too large for positional embedding grid")
        patches = patchify(t, P)
        ids_keep, ids_restore, seq_mask, len_keep = mask_sequence(noise, cfg["mask_ratio"])
        kept.append(patches.index_select(0, ids_keep))
        kept_pe.append(enc["pos_embedding"][:hp, :wp].reshape(-1, E).index_select(0, ids_keep))
        keep_lens.append(len_keep); full_lens.append(patches.shape[0])
        seq_masks.append(seq_mask); restores.append(ids_restore); dims.append((hp, wp))
    x = linear(pad_rows(kept), enc["projection.weight"], enc["projection.bias"]) + pad_rows(kept_pe)
    enc_mask = attention_mask(keep_lens, x.shape[1])
    dec_mask = attention_mask(full_lens, max(full_lens))
    x = encoder_stack(x, enc, "encoder_blocks", cfg["encoder_heads"], enc_mask)
    x = linear(x, sd["decoder_embed.weight"], sd["decoder_embed.bias"])
    Dd = x.shape[-1]
    rec = []
    for i in range(len(xs)):  # prepare_for_decoder, models.py:219-241
        seq = torch.cat([x[i, : keep_lens[i]], sd["mask_token"].reshape(1, Dd).repeat(full_lens[i] - keep_lens[i], 1)])
        seq = seq.index_select(0, restores[i])
        hp, wp = dims[i]
        rec.append(seq + sd["decoder_pos_embedding"][:hp, :wp].reshape(-1, Dd))
    h = encoder_stack(pad_rows(rec), {k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")},
                      "decoder_blocks", cfg["decoder_heads"], dec_mask)
    pred = linear(h, sd["decoder_unembed.weight"], sd["decoder_unembed.bias"])
    loss_mask = pad_rows([m.bool() for m in seq_masks], False) & ~dec_mask
    target = pad_rows([patchify(t, P) for t in ys])
    return pred, loss_mask, target


def mae_loss(pred, loss_mask, target):
    """MAELoss (models.py:271-288): per-patch normalised target (unbiased var, eps 1e-6), masked MSE."""
    mean = target.mean(dim=-1, keepdim=True)
    var = target.var(dim=-1, keepdim=True)  # unbiased, as torch.var defaults
    tn = (target - mean) / (var + 1.0e-6) ** 0.5
    loss = ((pred - tn) ** 2).mean(dim=-1)
    return (loss * loss_mask).sum() / loss_mask.sum()


# --------------------------------------------------------------------------- #
# scheduled sampling  (models.py:798-838)
# --------------------------------------------------------------------------- #
def gumbel_softmax(logits, gumbel, tau, hard):
    """F.gumbel_softmax with the Gumbel(0,1) draw supplied: softmax((logits + g) / tau); hard = straight-through one-hot."""
    y_soft = torch.softmax((logits.float() + gumbel) / tau, dim=-1)
    if not hard:
        return y_soft
    index = y_soft.max(dim=-1, keepdim=True)[1]
    y_hard = torch.zeros_like(y_soft).scatter_(-1, index, 1.0)
    return y_hard - y_soft.detach() + y_soft


def decoder_forward_embeddings(emb, memory, sd, H, lmx_pad_mask=None, mem_pad_mask=None):
    """OMRDecoder.forward(token_idxs_input=False) (models.py:463-483): embedding-space inputs + PE[:T]."""
    T = emb.shape[1]
    x = emb + sd["pos_embedding"][:T].unsqueeze(0)
    x = decoder_stack(x, memory, sd, H, True, lmx_pad_mask, mem_pad_mask)
    return linear(x, sd["unembed.weight"], sd["unembed.bias"])


def scheduled_sampling_forward(batch, sd, cfg, tf_prob, tau, hard, rand, gumbel):
    """ScheduledSamplingViTOMR.forward_train (models.py:819-834) with its two random draws supplied:
    `rand` (B,T) for the sampling mask (:803) and `gumbel` (B,T,V) for F.gumbel_softmax (:808)."""
    imgs, seqs = zip(*batch)
    enc_sd = {k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")}
    dec_sd = {k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}
    lat, lat_mask = encoder_forward(list(imgs), enc_sd, cfg["encoder"], fast_path_pad_fill=False)
    lat = transition_head(lat, sd)
    inp, tgt, lmx_mask = batchify_and_split_lmx_seqs(list(seqs), cfg["pad_idx"])
    logits = decoder_forward(inp, lat, dec_sd, cfg["decoder_heads"], lmx_mask, lat_mask)
    table = dec_sd["vocab_embedding.weight"]
    sample_mask = rand < (1 - tf_prob)
    gold = table[inp]                                                   # padding_idx row is zero in a reference state_dict
    expected = gumbel_softmax(logits, gumbel, tau, hard) @ table
    expected = torch.cat([gold[:, 0:1], expected], dim=1)[:, :-1]       # right shift (:813-815)
    mixed = torch.where(sample_mask.unsqueeze(-1), expected, gold)
    pred = decoder_forward_embeddings(mixed, lat, dec_sd, cfg["decoder_heads"], lmx_mask, lat_mask)
    return pred, tgt


# ------------------------------------------------------------------------------------------ GRPO
def topk_sample_step(logits, top_k, temperature, u):
    """One step of GRPOViTOMR.cached_forward_rollout_policy (models.py:1003-1029) for a (R, V) fp32 logits matrix with the
    multinomial draw written out: keep the top_k logits of each row (ties: lower index first), p = softmax(kept /
    temperature), the token is the first kept index (in vocabulary order) whose cumulative probability exceeds u[r]
    (inverse CDF - the same distribution torch.multinomial samples), log-prob = log_softmax(kept logits)[token] without
    the temperature.  Pure-Python loops: small cases only."""
    R, V = logits.shape
    toks, lps = [], []
    for r in range(R):
        row = [float(v) for v in logits[r]]
        order = sorted(range(V), key=lambda c: (-row[c], c))[: min(top_k, V)]
        kept = sorted(order)
        mx = max(row[c] for c in kept)
        w = [math.exp((row[c] - mx) / temperature) for c in kept]
        target, cum, choice = float(u[r]) * sum(w), 0.0, kept[-1]
        for c, wc in zip(kept, w):
            cum += wc
            if cum > target:
                choice = c
                break
        toks.append(choice)
        lps.append((row[choice] - mx) - math.log(sum(math.exp(row[c] - mx) for c in kept)))
    return torch.tensor(toks, dtype=torch.long), torch.tensor(lps, dtype=torch.float32)


def cached_forward_rollout_policy(memory, sd, H, uniforms, mem_pad_mask=None, max_actions=768, top_k=50, temperature=1.2,
                                  pad_idx=1, bos_idx=0, eos_idx=2):
    """GRPOViTOMR.cached_forward_rollout_policy (models.py:991-1049) with the draws supplied: uniforms[r, t] drives
    token t.  Same cached decoder step (and `time_step = t` quirk) as cached_greedy_generate above."""
    R = memory.shape[0]
    st = prepare_caches(memory, sd, H)
    rollouts = torch.full((R, max_actions), pad_idx, dtype=torch.long)
    rollouts[:, 0] = bos_idx
    logp = torch.zeros(R, max_actions, dtype=torch.float32)
    finished = torch.zeros(R, dtype=torch.bool)
    for t in range(1, max_actions):
        logits = cached_generate(rollouts[:, t - 1 : t], t, st, sd, H, mem_pad_mask).squeeze(1)
        tok, lp = topk_sample_step(logits.float(), top_k, temperature, uniforms[:, t])
        rollouts[:, t] = tok
        logp[:, t] = lp
        finished |= tok == eos_idx
        if bool(finished.all()):
            break
    m = create_inference_mask(rollouts, eos_idx)
    rollouts = rollouts.masked_fill(~m, pad_idx)
    n = int(m.sum(dim=-1).max())
    rollouts, logp, m = rollouts[:, :n], logp[:, :n], m[:, :n]
    return rollouts, logp.masked_fill(~m, 0.0), m


def grpo_objective(theta_logits, rollouts, rollout_attention_mask, old_policy_log_probs, advantages, epsilon, num_groups):
    """omr_grpo_train.py:240-262, statement by statement."""
    theta_log_probs = torch.log_softmax(theta_logits, dim=-1)
    theta_log_probs = torch.gather(theta_log_probs, dim=-1, index=rollouts[:, 1:].unsqueeze(-1)).squeeze(-1)
    ratios = torch.exp(theta_log_probs - old_policy_log_probs[:, 1:])
    unclipped = ratios * advantages.unsqueeze(1)
    clipped = torch.clip(ratios, min=1 - epsilon, max=1 + epsilon) * advantages.unsqueeze(1)
    unclipped = unclipped.masked_fill(rollout_attention_mask, 0)
    clipped = clipped.masked_fill(rollout_attention_mask, 0)
    lens = (~rollout_attention_mask).sum(dim=-1)
    obj = torch.sum(torch.min(unclipped, clipped), dim=-1) / lens
    return torch.sum(obj) / num_groups


def entropy_bonus(theta_logits, rollout_attention_mask, vocab_size):
    """omr_grpo_train.py:265-282."""
    probs = torch.softmax(theta_logits, dim=-1)
    ent = (-probs * torch.log_softmax(theta_logits, dim=-1)).sum(dim=-1).masked_fill(rollout_attention_mask, 0)
    ent = ent.sum(dim=-1) / (~rollout_attention_mask).sum(dim=-1)
    return ent.mean() / torch.log(torch.tensor(float(vocab_size)))


# ---------------------------------------------------------------------------------------------------------------------
# Input pipeline: v2.ToImage + v2.ToDtype(float32, scale=True) + DynamicResize (omr_teacher_force_train.py:278-282,
# pre_train.py:162-166, utils/utils.py:334-370).  The arithmetic is torchvision F.resize(BICUBIC, antialias=True) =
# torch's separable antialiased bicubic (third party: torch, aten/src/ATen/native/cpu/UpSampleKernel.cpp
# `_compute_indices_min_size_weights_aa` + `HelperInterpCubic::aa_filter`), restated here as two explicit weight matrices.
# Pinned by tests/golden/golden_resize.pt (outputs of the reference's own DynamicResize class).
def dynamic_resize_target(height, width, patch_size, max_seq_len):
    """utils/utils.py:346-354."""
    if width > height:
        aspect_ratio = width // height
        target_height = patch_size * math.floor(math.sqrt(max_seq_len / aspect_ratio))
        return target_height, target_height * aspect_ratio
    aspect_ratio = height // width
    target_width = patch_size * math.floor(math.sqrt(max_seq_len / aspect_ratio))
    return target_width * aspect_ratio, target_width


def _cubic_aa(x, a=-0.5):
    x = abs(x)
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1.0
    if x < 2.0:
        return (((x - 5.0) * x + 8.0) * x - 4.0) * a
    return 0.0


def aa_bicubic_matrix(in_size, out_size):
    """[out_size, in_size] fp32 matrix of normalised antialiased-bicubic weights (fp32 index arithmetic, as torch)."""
    f32 = lambda v: torch.tensor(v, dtype=torch.float32).item()  # noqa: E731  (round every intermediate to fp32)
    scale = f32(in_size / out_size)
    support = f32(2.0 * scale) if scale >= 1.0 else 2.0
    invscale = f32(1.0 / scale) if scale >= 1.0 else 1.0
    m = torch.zeros(out_size, in_size, dtype=torch.float32)
    for i in range(out_size):
        center = f32(scale * (i + 0.5))
        xmin = max(int(f32(center - support + 0.5)), 0)
        xsize = min(int(f32(center + support + 0.5)), in_size) - xmin
        w = [f32(_cubic_aa(f32(f32(j + xmin - center + 0.5) * invscale))) for j in range(xsize)]
        total = f32(sum(w))
        for j in range(xsize):
            m[i, xmin + j] = w[j] / total
    return m


def dynamic_resize(img, patch_size, max_seq_len, pe_max_height, pe_max_width, crop_imgs):
    """img: (1, H, W) uint8 or fp32 CPU tensor -> (1, h, w) fp32, utils/utils.py:343-370 statement by statement (the uint8
    case first applies v2.ToDtype(float32, scale=True): x / 255)."""
    x = img.to(torch.float32) / 255.0 if img.dtype == torch.uint8 else img.to(torch.float32)
    height, width = x.shape[-2], x.shape[-1]
    th, tw = dynamic_resize_target(height, width, patch_size, max_seq_len)
    x = x @ aa_bicubic_matrix(width, tw).t()                 # width first (fp32 intermediate) ...
    x = aa_bicubic_matrix(height, th) @ x                    # ... then height
    if crop_imgs:
        if th / patch_size > pe_max_height:
            ch = pe_max_height * patch_size
            top = int(round((th - ch) / 2.0))                # torchvision center_crop
            x = x[..., top:top + ch, :]
        if tw / patch_size > pe_max_width:
            cw = pe_max_width * patch_size
            left = int(round((tw - cw) / 2.0))
            x = x[..., :, left:left + cw]
    return x.clamp(0.0, 1.0)
