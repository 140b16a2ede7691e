"""Parameter shells + packed (cu_seqlens) transformer blocks over the sm_100a kernels.

The shells keep torch's parameter names (`self_attn.in_proj_weight`, `linear1.weight`,
`norm1.bias`, `layers.N`, `norm` ...) so reference checkpoints load unchanged, but
no torch compute module is ever called: every forward below goes through `acai_omr_b200.ops`.

Precision policy (mirrors the reference's use of autocast, vitomr_inference.py:80-84):
  * inside `torch.autocast("cuda", dtype=torch.bfloat16)`  -> bf16 operands / fp32 accumulate,
    weights cached as bf16 copies, LayerNorm statistics in fp32;
  * otherwise -> fp32 end to end (the token-exact mode).
Residual stream under bf16: torch.autocast keeps it in fp32 - LayerNorm autocasts to fp32 and a bf16 sub-layer output
added to an fp32 stream promotes - and only the Linear / attention operands are bf16.  Inference (no autograd) does the
same here (`stream_dtype`): every block carries the pair (x: fp32 stream rows, xb: the same rows in bf16 as the next GEMM's
operand), the out-projection / linear2 GEMMs read the fp32 residual and write fp32, LayerNorm writes both copies in one
pass.  Against the fp32 oracle the full-size encoder's error halves (l2 1.1e-2 -> 0.6e-2, profiles/r2a_parity.md) for
~10 % more time; ACAI_STREAM=bf16 selects round 1's single bf16 stream.  Training (autograd on) keeps the bf16 stream: it is
what the saved-for-backward tensors are.
"""
from __future__ import annotations

import math
import os
import weakref

import torch
from torch import nn

from . import functional as F
from . import ops


def compute_dtype() -> torch.dtype:
    if torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16:
        return torch.bfloat16
    return torch.float32


STREAM_FP32 = os.environ.get("ACAI_STREAM", "fp32") != "bf16"


def stream_dtype(dt: torch.dtype) -> torch.dtype:
    """dtype of the residual stream for compute dtype `dt` (see the module docstring)."""
    if dt == torch.bfloat16 and STREAM_FP32 and not torch.is_grad_enabled():
        return torch.float32
    return dt


def to_stream(h):
    """Packed rows `h` (compute dtype) -> (stream rows, GEMM operand rows)."""
    sdt = stream_dtype(h.dtype)
    return (h if sdt == h.dtype else h.to(sdt)), h


def norm_pair(norm, y, dt):
    """LayerNorm of stream rows y -> (stream rows, the same rows in the GEMM operand dtype `dt`), one kernel."""
    if y.dtype != dt:
        return norm(y, out_dtype=y.dtype, out2_dtype=dt)
    x = norm(y, out_dtype=dt)
    return x, x


def linear_residual_norm(a, lin, residual, norm, dt, dropout_p=0.0):
    """norm(residual + dropout(lin(a))) -> (stream rows, GEMM operand rows): the tail of a post-norm sub-layer.  Inference in
    bf16 runs it as ONE kernel (acai_linear_layernorm: the LayerNorm in the GEMM's epilogue across a thread-block cluster);
    otherwise the projection (+ residual in its epilogue) and the LayerNorm are two launches."""
    if (dropout_p == 0.0 and dt == torch.bfloat16 and a.dtype == dt and a.dim() == 2 and lin.bias is not None
            and not F.needs_grad(a, residual, lin.weight, lin.bias, norm.weight, norm.bias)):
        fused = ops.linear_layernorm(a, w_(lin.weight, dt), lin.bias, residual, norm.weight, norm.bias, norm.eps,
                                     want_f32=residual.dtype == torch.float32)
        if fused is not None:
            xb, x32 = fused
            return (xb, xb) if x32 is None else (x32, xb)
    y = F.linear(a, lin.weight, lin.bias, residual=residual, out_dtype=residual.dtype, dropout_p=dropout_p)
    return norm_pair(norm, y, dt)


def _can_fuse_tail(layer, x, dt):
    return (ops.FUSE_LINEAR_LN and dt == torch.bfloat16 and not (layer.training and layer.dropout_p > 0.0)
            and not torch.is_grad_enabled() and x.dim() == 2)


class _WeightCache:
    """bf16 copies of fp32 master parameters, refreshed when the parameter is updated in place."""

    def __init__(self):
        self._store = {}
        self._shadow = {}
        self.generation = 0

    def register_shadow(self, p: torch.Tensor, shadow: torch.Tensor):
        """`shadow` is a low-precision copy of `p` that its owner (train.FlatTrainer's fused AdamW) keeps in sync."""
        self._shadow[id(p)] = (p._version, p.data_ptr(), shadow, weakref.ref(p))

    def bump_generation(self):
        """Parameters were updated through raw pointers (no torch version bump): derived copies are stale."""
        self.generation += 1

    def get(self, p: torch.Tensor, dtype: torch.dtype, rows=None):
        if p is None:
            return None
        if p.dtype == dtype and rows is None:
            return p.detach()
        sh = self._shadow.get(id(p))
        if sh is not None and sh[3]() is not p:                   # the registered parameter died and its id was reused
            del self._shadow[id(p)]
            sh = None
        if sh is not None and sh[2].dtype == dtype:
            if sh[0] != p._version or sh[1] != p.data_ptr():     # written through torch since (load_state_dict, copy_)
                sh[2].copy_(p.detach())
                self._shadow[id(p)] = (p._version, p.data_ptr(), sh[2], sh[3])
            return sh[2] if rows is None else sh[2][rows[0]:rows[1]]
        key = (id(p), dtype, rows)
        hit = self._lookup(key, p)
        if hit is not None:
            return hit
        src = p.detach() if rows is None else p.detach()[rows[0]:rows[1]]
        return self._insert(key, p, src.to(dtype).contiguous())

    # An entry belongs to ONE parameter object: `id(p)` (and even its data pointer and version counter) come back once
    # that parameter has been freed, so every entry carries a weak reference that must still resolve to `p`, and dies
    # (releasing its device copy) with the parameter.
    def _lookup(self, key, p):
        hit = self._store.get(key)
        if hit is not None and hit[2]() is p and hit[0] == (p._version, p.data_ptr(), self.generation):
            return hit[1]
        return None

    def _insert(self, key, p, t):
        store = self._store
        self._store[key] = ((p._version, p.data_ptr(), self.generation), t, weakref.ref(p, lambda _r, k=key: store.pop(k, None)))
        return t

    def get_t(self, p: torch.Tensor, dtype: torch.dtype, rows=None):
        """Transposed copy [in, out] of (a row slice of) `p` in `dtype`: the B operand of dX = dY W."""
        key = (id(p), dtype, rows, "T")
        hit = self._lookup(key, p)
        if hit is not None:
            return hit
        return self._insert(key, p, ops.transpose_pad(self.get(p, dtype, rows), pad_to=1))

    def clear(self):
        self._store.clear()
        self._shadow.clear()


WEIGHTS = _WeightCache()


def w_(p, dtype, rows=None):
    """Parameter `p` (optionally a row slice) in the GEMM operand dtype."""
    return WEIGHTS.get(p, dtype, rows)


class Linear(nn.Linear):
    """nn.Linear parameters (same init); forward = acai_linear_fwd."""

    def forward(self, x, act=ops.ACT_NONE, residual=None, out_dtype=None):
        dt = compute_dtype()
        if x.dtype != dt:
            x = x.to(dt)
        return F.linear(x, self.weight, self.bias, act=act, residual=residual, out_dtype=out_dtype or dt)


class LayerNorm(nn.LayerNorm):
    """nn.LayerNorm parameters; forward = acai_layernorm_fwd (optionally fused residual add)."""

    def forward(self, x, residual=None, out_dtype=None, out2_dtype=None):
        return F.layernorm(x, self.weight, self.bias, self.eps, residual=residual,
                           out_dtype=out_dtype or x.dtype, out2_dtype=out2_dtype)


class MultiheadAttention(nn.Module):
    """Parameter layout and init of nn.MultiheadAttention(batch_first=True, bias=True):
    `in_proj_weight` [3D, D] = [Wq; Wk; Wv], `in_proj_bias` [3D], `out_proj.{weight,bias}`."""

    def __init__(self, embed_dim, num_heads, dropout=0.0, batch_first=True, bias=True, device=None, dtype=None):
        super().__init__()
        if embed_dim % num_heads != 0:
            raise AssertionError("embed_dim must be divisible by num_heads")
        if not (batch_first and bias):
            raise ValueError("only batch_first=True, bias=True are supported (what the reference uses)")
        kw = dict(device=device, dtype=dtype)
        self.embed_dim, self.num_heads, self.dropout, self.batch_first = embed_dim, num_heads, dropout, batch_first
        self.head_dim = embed_dim // num_heads
        self.in_proj_weight = nn.Parameter(torch.empty(3 * embed_dim, embed_dim, **kw))
        self.in_proj_bias = nn.Parameter(torch.empty(3 * embed_dim, **kw))
        self.out_proj = Linear(embed_dim, embed_dim, bias=True, **kw)
        nn.init.xavier_uniform_(self.in_proj_weight)
        nn.init.constant_(self.in_proj_bias, 0.0)
        nn.init.constant_(self.out_proj.bias, 0.0)

    # ---- packed helpers -------------------------------------------------------------------
    def _p(self):
        """Dropout probability in force: nn.MultiheadAttention(dropout=p) drops attention probabilities in training mode;
        the owning layer applies the same p to the block output before the residual add (dropout1/dropout2)."""
        return self.dropout if self.training else 0.0

    def _out(self, a, residual, dt, box=None):
        p = self._p()
        # residual + dropout(out_proj(a)): dropout and residual add both in the GEMM epilogue (acai_linear_dropout); the
        # sum comes out in the residual's dtype (fp32 when the stream is fp32)
        return F.linear(a, self.out_proj.weight, self.out_proj.bias, residual=residual,
                        out_dtype=dt if residual is None else residual.dtype, dropout_p=p, box_put=box)

    def self_attention_core(self, x, pk, causal, box=None):
        """softmax(q k^T) v over packed rows x:[T, D], heads concatenated, before the output projection.
        `box`: functional.GradBox of the block's residual add (training, residual rows = x)."""
        dt, D = x.dtype, self.embed_dim
        qkv = F.linear(x, self.in_proj_weight, self.in_proj_bias, out_dtype=dt, box_take=box)
        return F.self_attention(qkv, pk, self.num_heads, self.head_dim, causal,
                                flops=4.0 * D * pk.sum_sq * (0.5 if causal else 1.0), dropout_p=self._p())

    def self_attention_packed(self, x, pk, causal, residual):
        """out_proj(softmax(q k^T) v) + residual over packed rows x:[T, D] (x: GEMM operand rows, residual: stream rows)."""
        box = F.grad_box(x, residual)
        return self._out(self.self_attention_core(x, pk, causal, box), residual, x.dtype, box)

    def project_memory_packed(self, mem):
        """[K | V] = mem . W_kv^T + b_kv  (kv_caching.py:241-245) over packed memory rows."""
        dt, D = mem.dtype, self.embed_dim
        return F.linear(mem, self.in_proj_weight, self.in_proj_bias, out_dtype=dt, rows=(D, 3 * D))

    def cross_attention_core(self, x, kv, pq, pk, box=None):
        dt, D = x.dtype, self.embed_dim
        q = F.linear(x, self.in_proj_weight, self.in_proj_bias, out_dtype=dt, rows=(0, D), box_take=box)
        return F.cross_attention(q, kv, pq, pk, self.num_heads, self.head_dim,
                                 flops=4.0 * D * sum(a_ * b_ for a_, b_ in zip(pq.lens, pk.lens)), dropout_p=self._p())

    def cross_attention_packed(self, x, kv, pq, pk, residual):
        box = F.grad_box(x, residual)
        return self._out(self.cross_attention_core(x, kv, pq, pk, box), residual, x.dtype, box)


class Packed:
    """Ragged batch as packed rows: `cu` int32 device offsets, `lens` host list, `max_len`."""

    def __init__(self, lens, device):
        self.lens = [int(v) for v in lens]
        self.batch = len(self.lens)
        self.total = sum(self.lens)
        self.max_len = max(self.lens) if self.lens else 0
        self.sum_sq = float(sum(v * v for v in self.lens))
        off = [0]
        for v in self.lens:
            off.append(off[-1] + v)
        self.cu = ops.upload_small(torch.tensor(off, dtype=torch.int32), device)

    def mask(self, device, max_len=None):
        """True = padding, (B, L_max)  (create_attention_mask, models.py:70-73)."""
        L = max_len or self.max_len
        m = torch.arange(L)[None, :] >= torch.tensor(self.lens)[:, None]
        m = ops.upload_small(m, device)
        m._acai_lens = self.lens  # lets downstream modules re-pack without a device sync
        m._acai_lens_version = m._version
        return m


def attach_packed(t, packed):
    """Remember the packed rows a padded tensor was made from.  The note is honoured only while `t` has not been written
    to since (`Tensor._version`): after `t.mul_()`, `t[i] = ...` the consumer re-packs from the tensor's contents, as the
    reference - which only ever sees the padded tensor - would."""
    t._acai_packed = (t._version, packed)


def attached_packed(t):
    note = getattr(t, "_acai_packed", None)
    if note is None or note[0] != t._version:
        return None
    return note[1]


def lens_from_mask(mask, batch, length):
    """Host lengths from a (B, L) True=padding mask.  Uses the lengths attached by `Packed.mask`
    when present (and the mask has not been edited since); otherwise reads the mask back (one sync) and requires
    suffix-only padding."""
    if mask is None:
        return [length] * batch
    lens = getattr(mask, "_acai_lens", None)
    if lens is not None and getattr(mask, "_acai_lens_version", mask._version) != mask._version:
        lens = None                                          # written in place since: read the contents
    if lens is not None and len(lens) == batch and mask.shape[1] == length:
        return list(lens)
    m = mask.bool()
    lens_t = (~m).sum(dim=1)
    ok = (m == (torch.arange(length, device=m.device)[None, :] >= lens_t[:, None])).all()
    lens = lens_t.tolist()
    if not bool(ok):
        raise ValueError("acai_omr_b200 packs ragged batches: padding masks must mark a suffix of each row")
    return lens


class TransformerEncoderLayer(nn.Module):
    """Post-norm encoder layer with nn.TransformerEncoderLayer's parameter names (models.py:31)."""

    def __init__(self, d_model, nhead, dim_feedforward=2048, dropout=0.1, activation="gelu", layer_norm_eps=1e-5,
                 batch_first=True, norm_first=False, bias=True, device=None, dtype=None):
        super().__init__()
        if activation != "gelu" and activation is not torch.nn.functional.gelu:
            raise ValueError("only the exact-erf GELU activation is implemented")
        if norm_first:
            raise ValueError("only post-norm (norm_first=False) is implemented")
        kw = dict(device=device, dtype=dtype)
        self.self_attn = MultiheadAttention(d_model, nhead, dropout=dropout, **kw)
        self.linear1 = Linear(d_model, dim_feedforward, **kw)
        self.linear2 = Linear(dim_feedforward, d_model, **kw)
        self.norm1 = LayerNorm(d_model, eps=layer_norm_eps, **kw)
        self.norm2 = LayerNorm(d_model, eps=layer_norm_eps, **kw)
        self.dropout_p = dropout

    def forward_packed(self, x, pk: Packed, xb=None):
        """x: stream rows, xb: the same rows as GEMM operand (None: x) -> the pair after the layer (one tensor when xb is
        None, for callers that carry a single stream)."""
        single = xb is None
        xb = x if single else xb
        dt = xb.dtype
        if _can_fuse_tail(self, xb, dt):
            a = self.self_attn.self_attention_core(xb, pk, False)
            x, xb = linear_residual_norm(a, self.self_attn.out_proj, x, self.norm1, dt)
            h = F.linear(xb, self.linear1.weight, self.linear1.bias, act=ops.ACT_GELU, out_dtype=dt)
            x, xb = linear_residual_norm(h, self.linear2, x, self.norm2, dt)
            return x if single else (x, xb)
        y = self.self_attn.self_attention_packed(xb, pk, False, residual=x)
        x, xb = norm_pair(self.norm1, y, dt)
        x, xb = norm_pair(self.norm2, _mlp_residual(self, xb, dt, stream=x), dt)
        return x if single else (x, xb)


def _mlp_residual(layer, x, dt, stream=None):
    """x + dropout2(linear2(dropout(gelu(linear1(x)))))  (the feed-forward block of nn.Transformer*Layer, post-norm).
    `stream`: the residual rows when they are kept apart from the GEMM operand rows x (fp32 stream)."""
    p = layer.dropout_p if layer.training else 0.0
    l1, l2 = layer.linear1, layer.linear2
    if stream is not None and stream is not x:
        h = F.linear(x, l1.weight, l1.bias, act=ops.ACT_GELU, out_dtype=dt, dropout_p=p)
        return F.linear(h, l2.weight, l2.bias, residual=stream, out_dtype=stream.dtype, dropout_p=p)
    if not ops.FUSE_GELU_GEMM or l1.bias is None or l2.bias is None or x.dtype != dt:
        h = F.linear(x, l1.weight, l1.bias, act=ops.ACT_GELU, out_dtype=dt, dropout_p=p)
        return F.linear(h, l2.weight, l2.bias, residual=x, out_dtype=dt, dropout_p=p)
    if p > 0.0:
        box = F.grad_box(x, x)
        return F.dropout(F.feed_forward(x, l1.weight, l1.bias, l2.weight, l2.bias, add_residual=False, dropout_p=p, box_take=box),
                         p, True, residual=x, box_put=box)
    return F.feed_forward(x, l1.weight, l1.bias, l2.weight, l2.bias, add_residual=True)


def _check_dropout(module, p):
    """Dropout is implemented (counter-based masks, see functional.py); kept as a no-op hook for callers."""
    return None


class TransformerEncoder(nn.Module):
    """nn.TransformerEncoder-shaped container: `.layers`, `.norm`, `.num_layers` (models.py:30-34)."""

    def __init__(self, make_layer, num_layers, norm=None):
        super().__init__()
        self.layers = nn.ModuleList([make_layer() for _ in range(num_layers)])
        self.num_layers = num_layers
        self.norm = norm
        # nn.TransformerEncoder deep-copies one prototype layer: all layers start identical
        for layer in self.layers[1:]:
            layer.load_state_dict(self.layers[0].state_dict())

    def forward_packed(self, x, pk: Packed, final_dtype=None, xb=None, return_pair=False):
        """x: packed rows [sum L, D].  A caller that chains stacks passes the pair (x = stream rows, xb = operand rows) and
        asks for it back (`return_pair`); otherwise the pair is made and dropped here."""
        if xb is None:
            x, xb = to_stream(x)
        for layer in self.layers:
            _check_dropout(self, layer.dropout_p)
            x, xb = layer.forward_packed(x, pk, xb)
        if self.norm is not None:
            x = xb = self.norm(x, out_dtype=final_dtype or xb.dtype)
        return (x, xb) if return_pair else (x if self.norm is not None else xb)

    def forward(self, src, mask=None, src_key_padding_mask=None):
        """Padded API: src (B, L, D), src_key_padding_mask (B, L) True = padding."""
        if mask is not None:
            raise NotImplementedError("attention masks other than key padding are not used by the reference encoder")
        B, L, D = src.shape
        pk = Packed(lens_from_mask(src_key_padding_mask, B, L), src.device)
        dt = compute_dtype()
        x = F.pack_padded(src, pk.cu, pk.total, out_dtype=dt)
        x = self.forward_packed(x, pk, final_dtype=torch.float32 if self.norm is not None else None)
        fill = self.norm.bias if (self.norm is not None and not self.training) else None
        return F.unpack_padded(x, pk.cu, B, L, fill=fill)


class TransformerDecoderLayer(nn.Module):
    """Post-norm decoder layer with nn.TransformerDecoderLayer's parameter names (models.py:423)."""

    def __init__(self, d_model, nhead, dim_feedforward=2048, dropout=0.1, activation="gelu", layer_norm_eps=1e-5,
                 batch_first=True, norm_first=False, bias=True, device=None, dtype=None):
        super().__init__()
        if activation != "gelu" and activation is not torch.nn.functional.gelu:
            raise ValueError("only the exact-erf GELU activation is implemented")
        if norm_first:
            raise ValueError("only post-norm (norm_first=False) is implemented")
        kw = dict(device=device, dtype=dtype)
        self.self_attn = self._make_attention(d_model, nhead, dropout, **kw)
        self.multihead_attn = self._make_attention(d_model, nhead, dropout, **kw)
        self.linear1 = Linear(d_model, dim_feedforward, **kw)
        self.linear2 = Linear(dim_feedforward, d_model, **kw)
        self.norm1 = LayerNorm(d_model, eps=layer_norm_eps, **kw)
        self.norm2 = LayerNorm(d_model, eps=layer_norm_eps, **kw)
        self.norm3 = LayerNorm(d_model, eps=layer_norm_eps, **kw)
        self.dropout_p = dropout

    def _make_attention(self, d_model, nhead, dropout, **kw):
        return MultiheadAttention(d_model, nhead, dropout=dropout, **kw)

    def mlp_packed(self, x, xb=None):
        if xb is None:
            return self.norm3(_mlp_residual(self, x, x.dtype), out_dtype=x.dtype)
        return norm_pair(self.norm3, _mlp_residual(self, xb, xb.dtype, stream=x), xb.dtype)

    def forward_packed(self, x, tgt: Packed, mem, mem_pk: Packed, causal=True, xb=None):
        """x: stream rows, xb: operand rows (None: x; then one tensor comes back)."""
        single = xb is None
        xb = x if single else xb
        dt = xb.dtype
        if _can_fuse_tail(self, xb, dt):
            a = self.self_attn.self_attention_core(xb, tgt, causal)
            x, xb = linear_residual_norm(a, self.self_attn.out_proj, x, self.norm1, dt)
            kv = self.multihead_attn.project_memory_packed(mem)
            c = self.multihead_attn.cross_attention_core(xb, kv, tgt, mem_pk)
            x, xb = linear_residual_norm(c, self.multihead_attn.out_proj, x, self.norm2, dt)
            h = F.linear(xb, self.linear1.weight, self.linear1.bias, act=ops.ACT_GELU, out_dtype=dt)
            x, xb = linear_residual_norm(h, self.linear2, x, self.norm3, dt)
            return x if single else (x, xb)
        y = self.self_attn.self_attention_packed(xb, tgt, causal, residual=x)
        x, xb = norm_pair(self.norm1, y, dt)
        kv = self.multihead_attn.project_memory_packed(mem)
        y = self.multihead_attn.cross_attention_packed(xb, kv, tgt, mem_pk, residual=x)
        x, xb = norm_pair(self.norm2, y, dt)
        x, xb = self.mlp_packed(x, xb)
        return x if single else (x, xb)


class _Replayable:
    """fn for torch.utils.checkpoint: the first call (the forward) notes the dropout seed counter it starts from and lets it
    advance as usual; every later call (the re-run inside the backward) rewinds the counter to that value for its duration,
    so that it regenerates the first run's masks and leaves the seeds drawn around it alone."""

    def __init__(self, fn):
        self.fn, self.start = fn, None

    def __call__(self, *args):
        if self.start is None:
            self.start = ops.DropoutState.counter
            return self.fn(*args)
        saved = ops.DropoutState.counter
        ops.DropoutState.counter = self.start
        try:
            return self.fn(*args)
        finally:
            ops.DropoutState.counter = saved


class TransformerDecoder(nn.Module):
    """nn.TransformerDecoder-shaped container (models.py:422-426)."""

    def __init__(self, make_layer, num_layers, norm=None):
        super().__init__()
        self.layers = nn.ModuleList([make_layer() for _ in range(num_layers)])
        self.num_layers = num_layers
        self.norm = norm
        for layer in self.layers[1:]:
            layer.load_state_dict(self.layers[0].state_dict())

    def forward_packed(self, x, tgt: Packed, mem, mem_pk: Packed, causal=True, final_dtype=None, stream=None, checkpoint=False):
        """x: operand rows (compute dtype); `stream`: the same rows before they were rounded to it (fp32 embedding sum), if
        the caller has them.  `checkpoint` (OMRDecoder.forward(checkpoint_grads=True), models.py:470-478): per-layer
        activation checkpointing - only each layer's input is kept for the backward and the layer is run again there, with
        the same dropout masks (the seed counter is rewound to the value the first run started from)."""
        if checkpoint and torch.is_grad_enabled() and x.requires_grad:
            from torch.utils.checkpoint import checkpoint as ckpt
            for layer in self.layers:
                x = ckpt(_Replayable(layer.forward_packed), x, tgt, mem, mem_pk, causal, use_reentrant=False)
            if self.norm is not None:
                x = self.norm(x, out_dtype=final_dtype or x.dtype)
            return x
        if stream is not None and stream_dtype(x.dtype) == stream.dtype:
            xb = x
            x = stream
        else:
            x, xb = to_stream(x)
        for layer in self.layers:
            _check_dropout(self, layer.dropout_p)
            x, xb = layer.forward_packed(x, tgt, mem, mem_pk, causal, xb)
        if self.norm is not None:
            return self.norm(x, out_dtype=final_dtype or xb.dtype)
        return xb

    def forward(self, tgt, memory, tgt_mask=None, memory_mask=None, tgt_key_padding_mask=None,
                memory_key_padding_mask=None, tgt_is_causal=None, memory_is_causal=False):
        """Padded API of nn.TransformerDecoder.forward.  `tgt_mask` must be the causal triu mask (or None);
        it is not read: causality is applied inside the attention kernel."""
        if memory_mask is not None:
            raise NotImplementedError("memory_mask is not used by the reference")
        B, T, D = tgt.shape
        S = memory.shape[1]
        tp = Packed(lens_from_mask(tgt_key_padding_mask, B, T), tgt.device)
        mp = Packed(lens_from_mask(memory_key_padding_mask, B, S), tgt.device)
        dt = compute_dtype()
        x = F.pack_padded(tgt, tp.cu, tp.total, out_dtype=dt)
        mem = F.pack_padded(memory, mp.cu, mp.total, out_dtype=dt)
        x = self.forward_packed(x, tp, mem, mp, causal=tgt_mask is not None or bool(tgt_is_causal),
                                final_dtype=torch.float32 if self.norm is not None else None)
        return F.unpack_padded(x, tp.cu, B, T)


def gelu_mlp_head(x, lin_a: Linear, lin_b: Linear, out_dtype=None, dropout_p=0.0):
    """Linear -> exact GELU -> (Dropout) -> Linear with the activation fused into the first GEMM's epilogue."""
    dt = compute_dtype()
    if x.dtype != dt:
        x = x.to(dt)
    h = F.linear(x, lin_a.weight, lin_a.bias, act=ops.ACT_GELU, out_dtype=dt, dropout_p=dropout_p)
    return F.linear(h, lin_b.weight, lin_b.bias, out_dtype=out_dtype or dt)


SQRT_HALF = 1.0 / math.sqrt(2.0)
