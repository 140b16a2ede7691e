"""Differentiable front of `ops`: each function runs the forward kernel and, when a gradient is wanted,
records a `torch.autograd.Function` whose backward is again only sm_100a kernels (GEMMs for dX / dW on the
tcgen05 path, acai_layernorm_bwd, acai_attn_varlen_bwd, acai_ce_loss ...).  torch.autograd is used as the
tape and for gradient accumulation into `.grad`; it never computes a gradient itself.

Covers the training rows of SURVEY §8: teacher-forced ViTOMR (models.py:649-737 + OMRCELoss 784-796) and MAE
pre-training (models.py:197-288).  With gradients disabled every function is a plain `ops` call, so the
inference path is unchanged.

Dropout (training mode, p > 0) uses counter-based masks regenerated in the backward (`ops.DropoutState`): the keep
rate and 1/(1-p) scaling are nn.Dropout's, the random stream is not torch's Philox (it cannot be reproduced).

Not differentiable here (raise instead of silently detaching): interpolated position embeddings (oversize
images) and the second LayerNorm output used by the cached decoder.
"""
from __future__ import annotations

import os

import torch

from . import ops


def needs_grad(*tensors) -> bool:
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


def _weights():
    from .layers import WEIGHTS
    return WEIGHTS


def _slice(p, rows):
    return p if (p is None or rows is None) else p[rows[0]:rows[1]]


def _owned_grad(p):
    """The parameter's preallocated fp32 gradient buffer when a trainer owns it (train.FlatTrainer sets `_acai_grad_owned`):
    backward kernels then accumulate into it directly and report no gradient to autograd."""
    if p is not None and getattr(p, "_acai_grad_owned", False) and p.grad is not None and p.grad.dtype == torch.float32:
        return p.grad
    return None


def _full_rows(grad_rows, like, rows):
    """Gradient of a row slice -> gradient of the whole parameter."""
    if rows is None:
        return grad_rows
    full = torch.zeros(like.shape, dtype=grad_rows.dtype, device=grad_rows.device)
    full[rows[0]:rows[1]] = grad_rows
    return full


# ------------------------------------------------------------------------------------------ residual-gradient hand-over
class GradBox:
    """A post-norm sub-layer reads its input rows x twice: as the operand of its first GEMM and as the residual added to its
    output.  Autograd would add the two gradients of x with an elementwise kernel (71 `at::add` launches, 2.5 ms of the
    teacher-forced step).  Instead the node that owns the residual add (`put`: out_proj / the block's output dropout) leaves
    the residual gradient here and reports none, and the node of the first GEMM (`take`: in_proj / linear1), whose backward
    always runs later, adds it inside its dX GEMM's epilogue (`residual=`).  Only used when both uses are the SAME tensor."""
    __slots__ = ("grad",)

    def __init__(self):
        self.grad = None

    def take(self, like):
        g, self.grad = self.grad, None
        if g is not None and (g.dtype != like.dtype or g.numel() != like.numel()):
            g = g.to(like.dtype)
        return g


GRAD_BOX = os.environ.get("ACAI_GRAD_BOX", "1") != "0"


def grad_box(x, residual):
    """A GradBox when `residual` is the very tensor `x` and it needs a gradient (else None: the plain autograd path)."""
    return GradBox() if (GRAD_BOX and residual is x and x is not None and x.requires_grad and torch.is_grad_enabled()) else None


# ------------------------------------------------------------------------------------------ linear (+GELU, +residual)
class _Linear(torch.autograd.Function):
    """y = act(x W^T + b) + residual.   dX = dZ W,  dW = dZ^T X (fp32 out),  db = colsum(dZ),  dres = dY.
    In training the GELU is not fused into the GEMM epilogue: the pre-activation z is kept for the backward."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual, act, out_dtype, rows, dropout_p=0.0, dropout_seed=0, box_put=None, box_take=None):
        dt = x.dtype
        ctx.box_put, ctx.box_take = box_put, box_take
        K = x.shape[-1]
        x2 = x.reshape(-1, K)
        if x2.stride(-1) != 1:
            x2 = x2.contiguous()
        w = _weights().get(weight, dt, rows)
        b = _slice(bias, rows)
        b = b.detach() if b is not None else None
        z = None
        if act == ops.ACT_GELU:
            if residual is not None:
                raise ops.AcaiError("linear: GELU + residual is not used by the model")
            z = ops.linear(x2, w, b, out_dtype=dt)
            y = ops.gelu(z, out_dtype=out_dtype, dropout_p=dropout_p, dropout_seed=dropout_seed)
        elif dropout_p > 0.0:           # y = residual + dropout(x W^T + b): the post-norm blocks' nn.Dropout, in the GEMM epilogue
            y = _linear_out_dropout(x2, w, b, residual.reshape(-1, w.shape[0]).detach() if residual is not None else None,
                                    out_dtype, dropout_p, dropout_seed)
        else:
            y = ops.linear(x2, w, b, residual=residual.detach() if residual is not None else None, out_dtype=out_dtype)
        ctx.save_for_backward(x2, weight, z)
        ctx.rows, ctx.x_shape, ctx.has_bias, ctx.bias_ref = rows, x.shape, bias is not None, bias
        ctx.drop = (dropout_p, dropout_seed)
        ctx.res_meta = (residual.shape, residual.dtype) if residual is not None else None
        return y.view(*x.shape[:-1], y.shape[-1])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, z = ctx.saved_tensors
        dt, rows = x2.dtype, ctx.rows
        N = dy.shape[-1]
        dy2 = dy.reshape(-1, N)
        if z is not None:
            dz = ops.gelu_bwd(dy2, z, out_dtype=dt, dropout_p=ctx.drop[0], dropout_seed=ctx.drop[1])
        else:
            dz = dy2 if dy2.dtype == dt else dy2.to(dt)
            if not dz.is_contiguous():
                dz = dz.contiguous()
            if ctx.drop[0] > 0.0:        # through the output dropout (same counter-based mask)
                dz = ops.dropout(dz, ctx.drop[0], ctx.drop[1])
        dx = dw = db = dres = None
        if ctx.needs_input_grad[0]:
            # (the residual gradient of the block this GEMM opens, left by the node that closed it: GradBox)
            rg = ctx.box_take.take(x2) if ctx.box_take is not None else None
            rg = rg.reshape(x2.shape) if rg is not None else None
            # dX = dZ W straight from the forward's [out, in] weight copy (MN-major B operand); else from a transposed copy
            dx = ops.linear_dgrad(dz, _weights().get(weight, dt, rows), residual=rg) if dz.dim() == 2 else None
            if dx is None:
                dx = ops.linear(dz, _weights().get_t(weight, dt, rows), residual=rg, out_dtype=dt)
            dx = dx.view(ctx.x_shape)
        # the bias gradient rides in the weight-gradient kernel's pass over dz when both are wanted (ops.wgrad_accumulate)
        want_b = ctx.has_bias and ctx.needs_input_grad[2]
        db_target = db_full = None
        if want_b:
            g = _owned_grad(ctx.bias_ref)
            if g is not None:
                db_target = g if rows is None else g[rows[0]:rows[1]]
            else:
                db_full = torch.zeros(weight.shape[0], dtype=torch.float32, device=dz.device)
                db_target = db_full if rows is None else db_full[rows[0]:rows[1]]
                db = db_full
        if ctx.needs_input_grad[1]:
            g = _owned_grad(weight)
            if g is not None:            # accumulate straight into the (flat) gradient buffer: no temporary, no add kernel
                ops.wgrad_accumulate(dz, x2, g if rows is None else g[rows[0]:rows[1]], dbias=db_target)
            else:
                dw = torch.zeros(weight.shape if rows is None else (rows[1] - rows[0], weight.shape[1]), dtype=torch.float32,
                                 device=dz.device)
                dw = _full_rows(ops.wgrad_accumulate(dz, x2, dw, dbias=db_target), weight, rows)
        elif want_b:
            ops.colsum(dz, out=db_target, accumulate=True)
        if ctx.res_meta is not None and ctx.needs_input_grad[3]:
            shape, rdt = ctx.res_meta
            dres = dy.reshape(shape)
            if dres.dtype != rdt:
                dres = dres.to(rdt)
            if ctx.box_put is not None:
                ctx.box_put.grad, dres = dres, None
        return dx, dw, db, dres, None, None, None, None, None, None, None


def _linear_out_dropout(x2, w, b, residual, out_dtype, p, seed):
    y = ops.linear_dropout(x2, w, b, residual, p, seed) if out_dtype == torch.bfloat16 and x2.dtype == torch.bfloat16 else None
    if y is None:
        y = ops.dropout(ops.linear(x2, w, b, out_dtype=out_dtype), p, seed, residual=residual)
    return y


def linear(x, weight, bias=None, act=ops.ACT_NONE, residual=None, out_dtype=None, rows=None, dropout_p=0.0, box_put=None,
           box_take=None):
    """dropout(act(x @ weight[rows].T + bias[rows])) + residual; `weight` / `bias` are the (fp32 master) parameters.
    `box_put` / `box_take`: see GradBox (the residual gradient leaves through / the input gradient absorbs the box).
    `dropout_p` > 0: with the GELU activation the feed-forward block's inner nn.Dropout (fused into the GELU kernel); without
    an activation `residual + dropout(x W^T + b)`, the sub-layer output dropout (fused into the GEMM epilogue)."""
    dt = x.dtype
    seed = ops.DropoutState.next_seed() if dropout_p > 0.0 else 0
    if dropout_p > 0.0 and act != ops.ACT_GELU and not needs_grad(x, weight, bias, residual):
        x2 = x.reshape(-1, x.shape[-1])
        w = _weights().get(weight, dt, rows)
        y = _linear_out_dropout(x2 if x2.stride(-1) == 1 else x2.contiguous(), w, _slice(bias, rows),
                                residual.reshape(-1, w.shape[0]) if residual is not None else None, out_dtype or dt, dropout_p, seed)
        return y.view(*x.shape[:-1], y.shape[-1])
    if not needs_grad(x, weight, bias, residual):
        y = ops.linear(x, _weights().get(weight, dt, rows), _slice(bias, rows), act=act, residual=residual,
                       out_dtype=out_dtype or dt)
        return ops.dropout(y, dropout_p, seed) if dropout_p > 0.0 else y
    return _Linear.apply(x, weight, bias, residual, act, out_dtype or dt, rows, dropout_p, seed, box_put, box_take)


def _weight_grad(weight, dz, x2):
    """dW = dz^T x2: accumulated into the trainer-owned gradient (returns None) or returned as a fresh fp32 tensor."""
    g = _owned_grad(weight)
    if g is not None:
        ops.wgrad_accumulate(dz, x2, g)
        return None
    return ops.wgrad_accumulate(dz, x2, torch.zeros(weight.shape, dtype=torch.float32, device=dz.device))


def _bias_grad(bias, dz):
    g = _owned_grad(bias)
    if g is not None:
        ops.colsum(dz, out=g, accumulate=True)
        return None
    return ops.colsum(dz)


def _weight_bias_grad(weight, bias, dz, x2, need_w, need_b):
    """(dW, db) of one nn.Linear: when both are wanted, the bias gradient comes out of the weight-gradient kernel's own pass
    over dz (ops.wgrad_accumulate(dbias=...)); entries are None where the trainer owns the gradient buffer (or not needed)."""
    if not (need_w and need_b):
        return (_weight_grad(weight, dz, x2) if need_w else None), (_bias_grad(bias, dz) if need_b else None)
    gw, gb = _owned_grad(weight), _owned_grad(bias)
    dw = gw if gw is not None else torch.zeros(weight.shape, dtype=torch.float32, device=dz.device)
    db = gb if gb is not None else torch.zeros(bias.shape, dtype=torch.float32, device=dz.device)
    ops.wgrad_accumulate(dz, x2, dw, dbias=db)
    return (None if gw is not None else dw), (None if gb is not None else db)


class _FeedForward(torch.autograd.Function):
    """linear2(dropout(gelu(linear1(x)))) [+ x]: the feed-forward block as ONE tape node, so that the GELU (+ its dropout)
    rides in the up-projection GEMM's epilogue (acai_linear_gelu_train) and the GELU backward in the epilogue of the
    down-projection's dX GEMM (acai_linear_gelu_bwd); the block's residual gradient is fused into the last dX GEMM."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, add_residual, dropout_p, dropout_seed, box_take=None):
        dt = x.dtype
        ctx.box_take = box_take
        x2 = x.reshape(-1, x.shape[-1])
        if x2.stride(-1) != 1:
            x2 = x2.contiguous()
        W1, W2 = _weights().get(w1, dt, None), _weights().get(w2, dt, None)
        zh = ops.linear_gelu_train(x2, W1, b1.detach(), dropout_p, dropout_seed) if dt == torch.bfloat16 else None
        if zh is None:
            z = ops.linear(x2, W1, b1.detach(), out_dtype=dt)
            h = ops.gelu(z, out_dtype=dt, dropout_p=dropout_p, dropout_seed=dropout_seed)
        else:
            z, h = zh
        y = ops.linear(h, W2, b2.detach(), residual=x2 if add_residual else None, out_dtype=dt)
        ctx.save_for_backward(x2, z, h, w1, b1, w2, b2)
        ctx.meta = (x.shape, add_residual, dropout_p, dropout_seed)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, z, h, w1, b1, w2, b2 = ctx.saved_tensors
        shape, add_residual, p, seed = ctx.meta
        dt = x2.dtype
        dy2 = dy.reshape(-1, dy.shape[-1])
        if dy2.dtype != dt:
            dy2 = dy2.to(dt)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dz = ops.linear_dgrad(dy2, _weights().get(w2, dt, None), z=z, dropout_p=p, dropout_seed=seed) if ops.FUSE_GELU_GEMM else None
        if dz is None:
            w2t = _weights().get_t(w2, dt, None)                              # [hidden, d]
            dz = ops.linear_gelu_bwd(dy2, w2t, z, p, seed) if dt == torch.bfloat16 else None
            if dz is None:
                dz = ops.gelu_bwd(ops.linear(dy2, w2t, out_dtype=dt), z, out_dtype=dt, dropout_p=p, dropout_seed=seed)
        dx = dw1 = db1 = dw2 = db2 = None
        dw2, db2 = _weight_bias_grad(w2, b2, dy2, h, ctx.needs_input_grad[3], ctx.needs_input_grad[4])
        if ctx.needs_input_grad[0]:
            rg = dy2 if add_residual else (ctx.box_take.take(x2) if ctx.box_take is not None else None)
            rg = rg.reshape(x2.shape) if rg is not None else None
            dx = ops.linear_dgrad(dz, _weights().get(w1, dt, None), residual=rg)
            if dx is None:
                dx = ops.linear(dz, _weights().get_t(w1, dt, None), residual=rg, out_dtype=dt)
            dx = dx.view(shape)
        dw1, db1 = _weight_bias_grad(w1, b1, dz, x2, ctx.needs_input_grad[1], ctx.needs_input_grad[2])
        return dx, dw1, db1, dw2, db2, None, None, None, None


def feed_forward(x, w1, b1, w2, b2, add_residual=True, dropout_p=0.0, box_take=None):
    """linear2(dropout(gelu(linear1(x)))) (+ x): the feed-forward block of nn.Transformer*Layer up to (not including) its
    outer dropout; `dropout_p` is the inner nn.Dropout's rate.  `box_take`: GradBox of the residual add that follows."""
    dt = x.dtype
    seed = ops.DropoutState.next_seed() if dropout_p > 0.0 else 0
    if not needs_grad(x, w1, b1, w2, b2):
        h = linear(x, w1, b1, act=ops.ACT_GELU, out_dtype=dt, dropout_p=dropout_p)
        return linear(h, w2, b2, residual=x if add_residual else None, out_dtype=dt)
    return _FeedForward.apply(x, w1, b1, w2, b2, add_residual, dropout_p, seed, box_take)


# ------------------------------------------------------------------------------------------ LayerNorm
class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, gamma, beta, eps, out_dtype):
        y = ops.layernorm(x, gamma.detach(), beta.detach(), eps, residual=residual, out_dtype=out_dtype)
        ctx.save_for_backward(x, residual, gamma)
        ctx.eps, ctx.beta_ref = eps, beta
        return y

    @staticmethod
    def backward(ctx, dy):
        x, residual, gamma = ctx.saved_tensors
        gg, gb = _owned_grad(gamma), _owned_grad(ctx.beta_ref)
        if gg is not None and gb is not None and ctx.needs_input_grad[2] and ctx.needs_input_grad[3]:
            dv, dg, db = ops.layernorm_bwd(dy, x, residual, gamma.detach(), ctx.eps, dv_dtype=x.dtype, dgamma_acc=gg, dbeta_acc=gb)
        else:
            dv, dg, db = ops.layernorm_bwd(dy, x, residual, gamma.detach(), ctx.eps, dv_dtype=x.dtype)
        dres = None
        if residual is not None and ctx.needs_input_grad[1]:
            dres = dv if residual.dtype == dv.dtype else dv.to(residual.dtype)
        return dv, dres, dg, db, None, None


def layernorm(x, gamma, beta, eps, residual=None, out_dtype=None, out2_dtype=None):
    if not needs_grad(x, residual, gamma, beta):
        return ops.layernorm(x, gamma, beta, eps, residual=residual, out_dtype=out_dtype, out2_dtype=out2_dtype)
    if out2_dtype is not None:
        raise ops.AcaiError("layernorm: the two-output form is inference-only")
    return _LayerNorm.apply(x, residual, gamma, beta, eps, out_dtype or x.dtype)


# ------------------------------------------------------------------------------------------ attention
class _SelfAttention(torch.autograd.Function):
    """Packed ragged self-attention over a fused [T, 3D] QKV buffer; backward fills one d(QKV) buffer."""

    @staticmethod
    def forward(ctx, qkv, cu, max_len, heads, head_dim, causal, flops, dropout_p, seed):
        D = heads * head_dim
        out, lse = ops.attn_varlen(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], cu, cu, max_len, heads, head_dim,
                                   causal=causal, out_dtype=qkv.dtype, flops=flops, return_lse=True, dropout_p=dropout_p,
                                   dropout_seed=seed)
        ctx.save_for_backward(qkv, out, lse, cu)
        ctx.meta = (max_len, heads, head_dim, causal, dropout_p, seed)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse, cu = ctx.saved_tensors
        max_len, heads, head_dim, causal, dropout_p, seed = ctx.meta
        D = heads * head_dim
        dqkv = torch.empty_like(qkv)
        ops.attn_varlen_bwd(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], out, dout, lse, cu, cu, max_len, max_len, heads,
                            head_dim, causal, dqkv[:, :D], dqkv[:, D:2 * D], dqkv[:, 2 * D:], dropout_p=dropout_p,
                            dropout_seed=seed)
        return dqkv, None, None, None, None, None, None, None, None


class _CrossAttention(torch.autograd.Function):
    """q:[Tq, D] against a fused [Tk, 2D] K|V buffer (kv_caching.py:241-245 layout)."""

    @staticmethod
    def forward(ctx, q, kv, cu_q, cu_k, max_q, max_k, heads, head_dim, flops, dropout_p, seed):
        D = heads * head_dim
        out, lse = ops.attn_varlen(q, kv[:, :D], kv[:, D:], cu_q, cu_k, max_q, heads, head_dim, causal=False,
                                   out_dtype=q.dtype, flops=flops, return_lse=True, dropout_p=dropout_p, dropout_seed=seed)
        ctx.save_for_backward(q, kv, out, lse, cu_q, cu_k)
        ctx.meta = (max_q, max_k, heads, head_dim, dropout_p, seed)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, kv, out, lse, cu_q, cu_k = ctx.saved_tensors
        max_q, max_k, heads, head_dim, dropout_p, seed = ctx.meta
        D = heads * head_dim
        dq, dkv = torch.empty_like(q), torch.empty_like(kv)
        ops.attn_varlen_bwd(q, kv[:, :D], kv[:, D:], out, dout, lse, cu_q, cu_k, max_q, max_k, heads, head_dim, False,
                            dq, dkv[:, :D], dkv[:, D:], dropout_p=dropout_p, dropout_seed=seed)
        return dq, dkv, None, None, None, None, None, None, None, None, None


def self_attention(qkv, pk, heads, head_dim, causal, flops=0.0, dropout_p=0.0, dropout_seed=None):
    """`dropout_p` > 0: attention-probability dropout (training); a fresh seed is drawn unless one is given."""
    D = heads * head_dim
    seed = 0 if dropout_p <= 0.0 else (ops.DropoutState.next_seed() if dropout_seed is None else dropout_seed)
    if not needs_grad(qkv):
        return ops.attn_varlen(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], pk.cu, pk.cu, pk.max_len, heads, head_dim,
                               causal=causal, out_dtype=qkv.dtype, flops=flops, dropout_p=dropout_p, dropout_seed=seed)
    return _SelfAttention.apply(qkv, pk.cu, pk.max_len, heads, head_dim, causal, flops, dropout_p, seed)


def cross_attention(q, kv, pq, pk, heads, head_dim, flops=0.0, dropout_p=0.0, dropout_seed=None):
    D = heads * head_dim
    seed = 0 if dropout_p <= 0.0 else (ops.DropoutState.next_seed() if dropout_seed is None else dropout_seed)
    if not needs_grad(q, kv):
        return ops.attn_varlen(q, kv[:, :D], kv[:, D:], pq.cu, pk.cu, pq.max_len, heads, head_dim, causal=False,
                               out_dtype=q.dtype, flops=flops, dropout_p=dropout_p, dropout_seed=seed)
    return _CrossAttention.apply(q, kv, pq.cu, pk.cu, pq.max_len, pk.max_len, heads, head_dim, flops, dropout_p, seed)


# ------------------------------------------------------------------------------------------ packed <-> padded
class _Unpack(torch.autograd.Function):
    @staticmethod
    def forward(ctx, packed, cu, batch, max_len, fill, out_dtype):
        ctx.save_for_backward(cu)
        ctx.meta = (packed.shape[0], packed.dtype)
        return ops.unpack_padded(packed, cu, batch, max_len, fill=fill, out_dtype=out_dtype)

    @staticmethod
    def backward(ctx, dout):
        (cu,) = ctx.saved_tensors
        total, dt = ctx.meta
        return ops.pack_padded(dout, cu, total, out_dtype=dt), None, None, None, None, None


class _Pack(torch.autograd.Function):
    @staticmethod
    def forward(ctx, padded, cu, total, out_dtype):
        ctx.save_for_backward(cu)
        ctx.meta = (padded.shape[0], padded.shape[1], padded.dtype)
        return ops.pack_padded(padded, cu, total, out_dtype=out_dtype)

    @staticmethod
    def backward(ctx, dout):
        (cu,) = ctx.saved_tensors
        B, L, dt = ctx.meta
        return ops.unpack_padded(dout, cu, B, L, out_dtype=dt), None, None, None


def unpack_padded(packed, cu, batch, max_len, fill=None, out_dtype=None):
    """Padding rows are constants (`fill`) of the forward: they carry no gradient, as in the reference where
    they are masked out of every attention and of the loss."""
    if not needs_grad(packed):
        return ops.unpack_padded(packed, cu, batch, max_len, fill=fill, out_dtype=out_dtype)
    return _Unpack.apply(packed, cu, batch, max_len, fill.detach() if fill is not None else None, out_dtype)


def pack_padded(padded, cu, total, out_dtype=None):
    if not needs_grad(padded):
        return ops.pack_padded(padded, cu, total, out_dtype=out_dtype)
    return _Pack.apply(padded, cu, total, out_dtype)


# ------------------------------------------------------------------------------------------ token embedding
class _EmbedTokens(torch.autograd.Function):
    """table[ids] + pos[t] for (B, T) ids (models.py:466-470).  d(table) by row scatter-add (padding_idx rows get
    none, as nn.Embedding); d(pos[:T]) is the fixed-order sum over the batch."""

    @staticmethod
    def forward(ctx, ids, table, pos, out_dtype, padding_idx):
        B, T = ids.shape
        out = ops.embed_tokens(ids, table.detach(), pos.detach(), out_dtype, period=T)
        ctx.save_for_backward(ids)
        ctx.meta = (table.shape, pos.shape, padding_idx)
        return out

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        tshape, pshape, padding_idx = ctx.meta
        B, T = ids.shape
        dim = tshape[1]
        dtab = dpos = None
        if ctx.needs_input_grad[1]:
            idx = ids.reshape(-1).long()
            if padding_idx is not None:
                idx = torch.where(idx == padding_idx, torch.full_like(idx, -1), idx)
            dtab = ops.scatter_add_rows(dout.reshape(-1, dim), idx, torch.zeros(tshape, dtype=torch.float32, device=dout.device))
        if ctx.needs_input_grad[2]:
            dpos = torch.zeros(pshape, dtype=torch.float32, device=dout.device)
            ops.colsum(dout.reshape(B, T * dim), out=dpos.view(-1)[: T * dim])
        return None, dtab, dpos, None, None


def embed_tokens(ids, table, pos, out_dtype, padding_idx=None):
    if not needs_grad(table, pos):
        return ops.embed_tokens(ids, table, pos, out_dtype, period=ids.shape[1])
    return _EmbedTokens.apply(ids, table, pos, out_dtype, padding_idx)


# ------------------------------------------------------------------------------------------ patch embedding
class PatchGeometry:
    """Everything the patch-embedding kernels need about one ragged image batch (no arithmetic)."""

    def __init__(self, imgs, descs, pk, patch_size, channels, dims, patch_index=None):
        self.imgs, self.descs, self.pk, self.P, self.C, self.dims, self.patch_index = imgs, descs, pk, patch_size, channels, dims, patch_index

    def row_tables(self, pe_w, device):
        """Per output row: (row of the un-gathered patch matrix, row of the flattened position table)."""
        from .layers import Packed
        full_lens = [h * w for h, w in self.dims]
        full = Packed(full_lens, device)
        lens = ops.upload_small(torch.tensor(self.pk.lens), device)
        img = torch.repeat_interleave(torch.arange(len(self.dims), device=device), lens, output_size=self.pk.total)
        if self.patch_index is not None:
            p = self.patch_index.to(device=device, dtype=torch.int64)
        else:
            p = torch.arange(self.pk.total, device=device) - self.pk.cu.long()[img]
        wp = ops.upload_small(torch.tensor([w for _, w in self.dims]), device)[img]
        return full, full.cu.long()[img] + p, (p // wp) * pe_w + p % wp


class _PatchEmbed(torch.autograd.Function):
    """K1: rows = patches W^T + b + PE[row, col] (models.py:36-66).  Backward: dW = dY^T patches (GEMM),
    db = colsum(dY), dPE by row scatter-add.  Images carry no gradient."""

    @staticmethod
    def forward(ctx, weight, bias, pos_embedding, geom, out_dtype):
        out = ops.patch_embed(geom.imgs, geom.descs, geom.pk.cu, geom.pk.total, geom.P, geom.C, weight.detach(),
                              bias.detach() if bias is not None else None, pos_embedding.detach(), False, out_dtype,
                              patch_index=geom.patch_index)
        ctx.geom, ctx.shapes = geom, (weight.shape, pos_embedding.shape)
        return out

    @staticmethod
    def backward(ctx, dy):
        geom = ctx.geom
        wshape, pshape = ctx.shapes
        dev, dt = dy.device, dy.dtype
        dy = dy.contiguous()
        dw = db = dpe = None
        full, patch_row, pe_row = geom.row_tables(pshape[1], dev)
        if ctx.needs_input_grad[0]:
            patches = ops.patchify(geom.imgs, geom.descs, full.cu, full.total, geom.P, geom.C, out_dtype=dt)
            if geom.patch_index is not None:
                patches = ops.gather_rows(patches, patch_row)
            dw = ops.linear(ops.transpose_pad(dy), ops.transpose_pad(patches), out_dtype=torch.float32)
        if ctx.needs_input_grad[1]:
            db = ops.colsum(dy)
        if ctx.needs_input_grad[2]:
            dpe = torch.zeros(pshape, dtype=torch.float32, device=dev)
            ops.scatter_add_rows(dy, pe_row, dpe.view(-1, pshape[-1]))
        return dw, db, dpe, None, None


def patch_embed(geom: PatchGeometry, weight, bias, pos_embedding, interpolate, out_dtype):
    if not needs_grad(weight, bias, pos_embedding):
        return ops.patch_embed(geom.imgs, geom.descs, geom.pk.cu, geom.pk.total, geom.P, geom.C, weight, bias, pos_embedding,
                               interpolate, out_dtype, patch_index=geom.patch_index)
    if interpolate:
        raise ops.AcaiError("patch_embed: training through an interpolated position grid (oversize image) is not implemented")
    return _PatchEmbed.apply(weight, bias, pos_embedding, geom, out_dtype)


# ------------------------------------------------------------------------------------------ MAE restore
class _MaeRestore(torch.autograd.Function):
    """MAE.prepare_for_decoder on packed rows (models.py:219-241): out[b, i] = (r < keep_b ? latent[b, r] : mask_token)
    + dec_PE[i // w, i % w] with r = ids_restore[b][i]."""

    @staticmethod
    def forward(ctx, latent, mask_token, dec_pe, cu_keep, cu_full, ids_restore, wps, total_rows, out_dtype):
        out = ops.mae_restore(latent, cu_keep, cu_full, ids_restore, wps, total_rows, mask_token.detach(), dec_pe.detach(),
                              out_dtype)
        ctx.save_for_backward(cu_keep, cu_full, ids_restore, wps)
        ctx.meta = (latent.shape, latent.dtype, mask_token.shape, dec_pe.shape)
        return out

    @staticmethod
    def backward(ctx, dout):
        cu_keep, cu_full, ids_restore, wps = ctx.saved_tensors
        lshape, ldt, mshape, pshape = ctx.meta
        dev = dout.device
        dout = dout.contiguous()
        total = dout.shape[0]
        cuk, cuf = cu_keep.long(), cu_full.long()
        img = torch.repeat_interleave(torch.arange(cuf.numel() - 1, device=dev), cuf[1:] - cuf[:-1], output_size=total)
        i = torch.arange(total, device=dev) - cuf[img]
        r = ids_restore.long()
        kept = r < (cuk[1:] - cuk[:-1])[img]
        src = torch.full((lshape[0] + 1,), -1, dtype=torch.int64, device=dev)      # last slot swallows the masked rows
        src[torch.where(kept, cuk[img] + r, torch.full_like(r, lshape[0]))] = torch.arange(total, device=dev)
        src = src[:-1]
        dlat = dmask = dpe = None
        d_kept = ops.gather_rows(dout, src, out_dtype=ldt)                  # also the kept share of sum(dout)
        if ctx.needs_input_grad[0]:
            dlat = d_kept
        if ctx.needs_input_grad[1]:
            dmask = (ops.colsum(dout) - ops.colsum(d_kept)).view(mshape)
        if ctx.needs_input_grad[2]:
            wp = wps.long()[img]
            dpe = torch.zeros(pshape, dtype=torch.float32, device=dev)
            ops.scatter_add_rows(dout, (i // wp) * pshape[1] + i % wp, dpe.view(-1, pshape[-1]))
        return dlat, dmask, dpe, None, None, None, None, None, None


def mae_restore(latent, cu_keep, cu_full, ids_restore, wps, total_rows, mask_token, dec_pe, out_dtype):
    if not needs_grad(latent, mask_token, dec_pe):
        return ops.mae_restore(latent, cu_keep, cu_full, ids_restore, wps, total_rows, mask_token, dec_pe, out_dtype)
    return _MaeRestore.apply(latent, mask_token, dec_pe, cu_keep, cu_full, ids_restore, wps, total_rows, out_dtype)


# ------------------------------------------------------------------------------------------ losses
class _CELoss(torch.autograd.Function):
    """OMRCELoss (models.py:784-796): mean over non-pad targets of the (label-smoothed) cross entropy."""

    @staticmethod
    def forward(ctx, logits, targets, ignore_index, normaliser, label_smoothing):
        sums = ops.ce_loss(logits, targets, ignore_index, label_smoothing=label_smoothing)
        denom = sums[1] if normaliser is None else normaliser
        ctx.save_for_backward(logits, targets, denom)
        ctx.ignore_index, ctx.label_smoothing = ignore_index, label_smoothing
        ctx.mark_non_differentiable(sums)
        return sums[0] / denom, sums

    @staticmethod
    def backward(ctx, g, _gs):
        logits, targets, denom = ctx.saved_tensors
        scale = (g.float() / denom).reshape(1).contiguous()
        return (ops.ce_loss(logits, targets, ctx.ignore_index, dlogits_scale=scale, dlogits_dtype=logits.dtype,
                            label_smoothing=ctx.label_smoothing), None, None, None, None)


def ce_loss(logits, targets, ignore_index, normaliser=None, label_smoothing=0.0):
    """-> (loss, sums = [sum of token losses, token count]).  `normaliser` (1-element device tensor) replaces the local
    count, e.g. the global token count divided by the world size under data parallelism."""
    if not needs_grad(logits):
        sums = ops.ce_loss(logits, targets, ignore_index, label_smoothing=label_smoothing)
        return sums[0] / (sums[1] if normaliser is None else normaliser), sums
    return _CELoss.apply(logits, targets, ignore_index, normaliser, float(label_smoothing))


class _MAELoss(torch.autograd.Function):
    """MAELoss (models.py:271-288): mean over masked patches of mean_f (pred - normalised target)^2."""

    @staticmethod
    def forward(ctx, pred, loss_mask, target, normaliser):
        sums, _ = ops.mae_loss(pred, loss_mask, target)
        denom = sums[1] if normaliser is None else normaliser
        ctx.save_for_backward(pred, loss_mask, target, denom)
        ctx.mark_non_differentiable(sums)
        return sums[0] / denom, sums

    @staticmethod
    def backward(ctx, g, _gs):
        pred, loss_mask, target, denom = ctx.saved_tensors
        scale = (g.float() / denom).reshape(1).contiguous()
        return ops.mae_loss_bwd(pred, loss_mask, target, scale, dpred_dtype=pred.dtype), None, None, None


def mae_loss(pred, loss_mask, target, normaliser=None):
    if not needs_grad(pred):
        sums, _ = ops.mae_loss(pred, loss_mask, target)
        return sums[0] / (sums[1] if normaliser is None else normaliser), sums
    return _MAELoss.apply(pred, loss_mask, target, normaliser)


# ------------------------------------------------------------------------------------------ scheduled sampling (models.py:798-838)
class _GumbelSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, noise, tau, hard):
        y, y_soft = ops.gumbel_softmax(logits, noise, tau, hard)
        ctx.save_for_backward(y_soft)
        ctx.meta = (tau, logits.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y_soft,) = ctx.saved_tensors
        tau, dt = ctx.meta
        return ops.gumbel_softmax_bwd(dy, y_soft, tau, dt), None, None, None


def gumbel_softmax(logits, noise, tau, hard):
    """F.gumbel_softmax with the Gumbel noise supplied (straight-through gradient when `hard`)."""
    if not needs_grad(logits):
        return ops.gumbel_softmax(logits, noise, tau, hard)[0]
    return _GumbelSoftmax.apply(logits, noise, tau, hard)


class _ExpectedEmbedding(torch.autograd.Function):
    """probs[rows, V] @ table[V, D]  (models.py:811).  d probs = dout table^T,  d table += probs^T dout (wgrad GEMM)."""

    @staticmethod
    def forward(ctx, probs, table):
        dt = probs.dtype
        V = probs.shape[-1]
        p2 = probs.reshape(-1, V)
        out = ops.linear(p2, _weights().get_t(table, dt), out_dtype=dt)          # weight [D, V] = table^T
        ctx.save_for_backward(p2, table)
        ctx.shape = probs.shape
        return out.view(*probs.shape[:-1], table.shape[1])

    @staticmethod
    def backward(ctx, dout):
        p2, table = ctx.saved_tensors
        dt = p2.dtype
        d2 = dout.reshape(-1, table.shape[1])
        d2 = (d2 if d2.dtype == dt else d2.to(dt)).contiguous()
        dprobs = dtab = None
        if ctx.needs_input_grad[0]:
            dprobs = ops.linear(d2, _weights().get(table, dt), out_dtype=dt).view(ctx.shape)     # [rows, D] x [V, D]^T
        if ctx.needs_input_grad[1]:
            g = _owned_grad(table)
            if g is not None:
                ops.wgrad_accumulate(p2, d2, g)
            else:
                dtab = ops.wgrad_accumulate(p2, d2, torch.zeros(table.shape, dtype=torch.float32, device=d2.device))
        return dprobs, dtab


def expected_embedding(probs, table):
    if not needs_grad(probs, table):
        return ops.linear(probs, _weights().get_t(table, probs.dtype), out_dtype=probs.dtype)
    return _ExpectedEmbedding.apply(probs, table)


class _MixRows(torch.autograd.Function):
    """out[r] = (sel[r] >= 0 ? a[sel[r]] : b[r]) + pos[r % period]; every row of `a` is selected at most once."""

    @staticmethod
    def forward(ctx, b, a, sel, pos, period):
        out = ops.mix_rows(b, a, sel, pos.detach() if pos is not None else None, period)
        ctx.save_for_backward(sel)
        ctx.meta = (b.shape, b.dtype, a.shape if a is not None else None, a.dtype if a is not None else None,
                    pos.shape if pos is not None else None, period)
        return out

    @staticmethod
    def backward(ctx, dout):
        (sel,) = ctx.saved_tensors
        bshape, bdt, ashape, adt, pshape, period = ctx.meta
        dim = bshape[-1]
        d2 = dout.reshape(-1, dim).contiguous()
        rows = d2.shape[0]
        db = da = dpos = None
        ar = torch.arange(rows, device=d2.device)
        if ctx.needs_input_grad[0]:
            db = (d2 if sel is None else ops.gather_rows(d2, torch.where(sel.reshape(-1) < 0, ar, torch.full_like(ar, -1)))).view(bshape)
            if db.dtype != bdt:
                db = db.to(bdt)
        if ashape is not None and ctx.needs_input_grad[1]:
            n_a = 1
            for v in ashape[:-1]:
                n_a *= v
            inv = torch.full((n_a + 1,), -1, dtype=torch.int64, device=d2.device)
            s = sel.reshape(-1)
            inv[torch.where(s >= 0, s, torch.full_like(s, n_a))] = ar
            da = ops.gather_rows(d2, inv[:-1], out_dtype=adt).view(ashape)
        if pshape is not None and ctx.needs_input_grad[3]:
            dpos = torch.zeros(pshape, dtype=torch.float32, device=d2.device)
            ops.colsum(d2.view(rows // period, period * dim), out=dpos.view(-1)[: period * dim])
        return db, da, None, dpos, None


def mix_rows(b, a=None, sel=None, pos=None, period=0):
    if not needs_grad(b, a, pos):
        return ops.mix_rows(b, a, sel, pos, period)
    return _MixRows.apply(b, a, sel, pos, period)


class _EmbedRows(torch.autograd.Function):
    """table[ids] without positions (nn.Embedding with padding_idx: that row receives no gradient)."""

    @staticmethod
    def forward(ctx, ids, table, out_dtype, padding_idx):
        ctx.save_for_backward(ids)
        ctx.meta = (table.shape, padding_idx)
        return ops.gather_rows(table.detach(), ids, out_dtype=out_dtype).view(*ids.shape, table.shape[1])

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        tshape, padding_idx = ctx.meta
        idx = ids.reshape(-1).long()
        if padding_idx is not None:
            idx = torch.where(idx == padding_idx, torch.full_like(idx, -1), idx)
        dtab = ops.scatter_add_rows(dout.reshape(-1, tshape[1]), idx, torch.zeros(tshape, dtype=torch.float32, device=dout.device))
        return None, dtab, None, None


def embed_rows(ids, table, out_dtype, padding_idx=None):
    if not needs_grad(table):
        return ops.gather_rows(table, ids, out_dtype=out_dtype).view(*ids.shape, table.shape[1])
    return _EmbedRows.apply(ids, table, out_dtype, padding_idx)


# ------------------------------------------------------------------------------------------ dropout
class _Dropout(torch.autograd.Function):
    """y = residual + dropout(x): nn.Dropout followed by the residual add of a post-norm block."""

    @staticmethod
    def forward(ctx, x, residual, p, seed, box_put=None):
        ctx.meta = (p, seed, residual is not None and residual.dtype)
        ctx.box_put = box_put
        return ops.dropout(x, p, seed, residual=residual)

    @staticmethod
    def backward(ctx, dy):
        p, seed, rdt = ctx.meta
        dx = ops.dropout(dy, p, seed) if ctx.needs_input_grad[0] else None
        dres = None
        if rdt is not False and ctx.needs_input_grad[1]:
            dres = dy if dy.dtype == rdt else dy.to(rdt)
            if ctx.box_put is not None:
                ctx.box_put.grad, dres = dres, None
        return dx, dres, None, None, None


def dropout(x, p, training, residual=None, box_put=None):
    """nn.Dropout(p)(x) (+ residual).  Identity (or a plain add folded elsewhere by the caller) when not training or p == 0."""
    if not training or p <= 0.0:
        raise ops.AcaiError("functional.dropout: call only when a mask is needed (training and p > 0)")
    seed = ops.DropoutState.next_seed()
    if not needs_grad(x, residual):
        return ops.dropout(x, p, seed, residual=residual)
    return _Dropout.apply(x, residual, p, seed, box_put)
