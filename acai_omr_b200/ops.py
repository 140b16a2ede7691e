"""Tensor-level wrappers over the C ABI (include/acai_b200.h).

torch is used for device memory and the current stream only; every arithmetic op here is a
launch of one of this repo's sm_100a kernels.  CPU tensors are rejected: there is no fallback.
`launch_count()` reports how many kernels were enqueued through this module (bench.py's
`gpu_launches`).
"""
from __future__ import annotations

import ctypes as C
import os
import math

import torch

from . import _lib
from ._lib import AcaiError, ImageDesc

F32, BF16 = 0, 1
ACT_NONE, ACT_GELU = 0, 1
_DT = {torch.float32: F32, torch.bfloat16: BF16}
_launches = 0


def launch_count() -> int:
    return _launches


def _bump(n=1):
    global _launches
    _launches += n


class Profile:
    """Per-launch CUDA-event timing (bench.py's roofline leg).  Inside `with Profile() as p:` every kernel launched
    through this module is bracketed by two events on the launching stream; `p.summary()` groups device time,
    algorithmic FLOPs and algorithmic bytes by entry point."""
    active = None

    def __init__(self):
        self.records = []

    def __enter__(self):
        Profile.active = self
        return self

    def __exit__(self, *exc):
        Profile.active = None

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for name, e0, e1, flops, nbytes in self.records:
            d = out.setdefault(name, dict(launches=0, ms=0.0, flops=0.0, bytes=0.0))
            d["launches"] += 1
            d["ms"] += e0.elapsed_time(e1)
            d["flops"] += flops
            d["bytes"] += nbytes
        return out


# ACAI_NVTX=1: every C-ABI call sits in an NVTX range named after its entry point (the kernel family), so that
# `ncu --nvtx --nvtx-include "acai_attn_varlen_fwd/"` (or any NVTX-aware tool) can pick launches by family.  SURVEY section 5.
NVTX = os.environ.get("ACAI_NVTX", "0") == "1"


def _call(name, fn, *args, flops=0.0, nbytes=0.0):
    prof = Profile.active
    if prof is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    if NVTX:
        torch.cuda.nvtx.range_push(name)
        try:
            _lib.check(fn(*args), name)
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        _lib.check(fn(*args), name)
    if prof is not None:
        e1.record()
        prof.records.append((name, e0, e1, flops, nbytes))
    _bump()


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise AcaiError(f"unsupported dtype {t.dtype} (fp32 and bf16 only)") from None


def _cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise AcaiError("acai_omr_b200 ops need CUDA tensors (got a CPU tensor); there is no CPU fallback")
    _lib.require_device()


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    # (torch.cuda.current_stream() builds a Stream object: ~16 us per call, a third of an op's host cost)
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))


def _f32c(t):
    """fp32 contiguous view/copy of a (small) parameter tensor."""
    if t is None:
        return None
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _rows2d(x):
    """View x as [rows, cols] with a unit inner stride; returns (tensor, rows, cols, ld)."""
    cols = x.shape[-1]
    if x.dim() == 2 and x.stride(1) == 1:
        return x, x.shape[0], cols, x.stride(0)
    x2 = x.reshape(-1, cols)
    if x2.stride(1) != 1:
        x2 = x2.contiguous()
    return x2, x2.shape[0], cols, x2.stride(0)


# ------------------------------------------------------------------------------------------
def linear(x, weight, bias=None, act=ACT_NONE, residual=None, out_dtype=None, out=None):
    """act(x @ weight.T + bias) + residual over the last dim (acai_linear_fwd)."""
    _cuda(x, weight, bias, residual)
    x2, M, K, lda = _rows2d(x)
    w = weight.detach()
    if w.stride(-1) != 1:
        w = w.contiguous()
    N = w.shape[0]
    if w.shape[1] != K:
        raise AcaiError(f"linear: x has {K} features, weight expects {w.shape[1]}")
    if out is not None:
        if out.shape != (M, N) or out.stride(1) != 1:
            raise AcaiError("linear: `out` must be [M, N] with a unit inner stride")
        out_dtype = out.dtype
    else:
        out_dtype = out_dtype or x.dtype
        out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    if M == 0:
        return out.view(*x.shape[:-1], N)
    r2, ldr, rdt = None, 0, 0
    if residual is not None:
        r2, rm, rn, ldr = _rows2d(residual)
        if rm != M or rn != N:
            raise AcaiError("linear: residual shape mismatch")
        rdt = _dt(r2)
    b = _f32c(bias)
    lib = _lib.load()
    # A long bf16 GEMM whose contraction is not a multiple of 64 (the 227-class head's dX in training: [tokens, 227] x [1024, 227],
    # 0.7 ms per step on the SIMT kernel) runs on the tensor cores behind two zero-padded operand copies (~20 MB).
    if K % 64 != 0 and K > 64 and M >= 1024 and N >= 64 and x2.dtype == torch.bfloat16 and w.dtype == torch.bfloat16:
        Kp = (K + 63) // 64 * 64
        xp = torch.zeros((M, Kp), dtype=torch.bfloat16, device=x2.device)
        xp[:, :K].copy_(x2)
        wp = torch.zeros((N, Kp), dtype=torch.bfloat16, device=x2.device)
        wp[:, :K].copy_(w)
        x2, w, K, lda = xp, wp, Kp, Kp
    name = "acai_linear_fwd" if Profile.active is None else \
        f"acai_linear_fwd[{M}x{N}x{K}{',gelu' if act else ''}{',res' if r2 is not None else ''}]"
    _call(name, lib.acai_linear_fwd, _p(x2), _dt(x2), lda, _p(w), _dt(w), w.stride(0), _p(b), _p(r2), rdt, ldr,
                                   _p(out), _DT[out_dtype], out.stride(0), M, N, K, act, _stream(),
          flops=2.0 * M * N * K, nbytes=float(x2.element_size() * M * K + w.element_size() * N * K + out.element_size() * M * N))
    return out.view(*x.shape[:-1], N)


def layernorm(x, gamma, beta, eps, residual=None, out_dtype=None, out2_dtype=None):
    """LayerNorm(x + residual) (acai_layernorm_fwd).  Returns out, or (out, out2) if out2_dtype."""
    _cuda(x, gamma, beta, residual)
    x2, rows, cols, ld = _rows2d(x)
    if ld != cols:
        x2 = x2.contiguous()
    r2 = None
    if residual is not None:
        r2 = residual.reshape(-1, cols).contiguous()
        if r2.shape[0] != rows:
            raise AcaiError("layernorm: residual shape mismatch")
    out_dtype = out_dtype or x.dtype
    out = torch.empty((rows, cols), dtype=out_dtype, device=x.device)
    out2 = torch.empty((rows, cols), dtype=out2_dtype, device=x.device) if out2_dtype is not None else None
    if rows == 0:
        return out.view(x.shape) if out2 is None else (out.view(x.shape), out2.view(x.shape))
    g, b = _f32c(gamma), _f32c(beta)
    lib = _lib.load()
    _call("acai_layernorm_fwd", lib.acai_layernorm_fwd, _p(x2), _dt(x2), _p(r2), _dt(r2) if r2 is not None else 0, _p(g), _p(b),
                                      float(eps), _p(out), _DT[out_dtype], _p(out2),
                                      _DT[out2_dtype] if out2_dtype is not None else 0, rows, cols, _stream())
    out = out.view(x.shape)
    return out if out2 is None else (out, out2.view(x.shape))


def attn_varlen(q, k, v, cu_q, cu_k, max_q, heads, head_dim, causal=False, out_dtype=None, scale=None, flops=0.0,
                return_lse=False, dropout_p=0.0, dropout_seed=0):
    """Packed ragged attention.  q:[Tq, >=H*d] k,v:[Tk, >=H*d] row-strided views (acai_attn_varlen_fwd)."""
    _cuda(q, k, v, cu_q, cu_k)
    if not (q.dtype == k.dtype == v.dtype):
        raise AcaiError("attn_varlen: q/k/v dtype mismatch")
    for t in (q, k, v):
        if t.dim() != 2 or t.stride(1) != 1:
            raise AcaiError("attn_varlen: q/k/v must be 2-D with unit inner stride")
    out_dtype = out_dtype or q.dtype
    D = heads * head_dim
    out = torch.empty((q.shape[0], D), dtype=out_dtype, device=q.device)
    lse = torch.empty((heads, q.shape[0]), dtype=torch.float32, device=q.device) if return_lse else None
    lib = _lib.load()
    _call("acai_attn_varlen_fwd", lib.acai_attn_varlen_fwd, _p(q), _p(k), _p(v), _dt(q), q.stride(0), k.stride(0), v.stride(0),
                                        _p(cu_q), _p(cu_k), cu_q.numel() - 1, int(max_q), q.shape[0], k.shape[0], heads, head_dim,
                                        1 if causal else 0, float(scale or 1.0 / math.sqrt(head_dim)),
                                        _p(out), _DT[out_dtype], D, _p(lse), float(dropout_p), C.c_uint64(dropout_seed),
                                        _stream(), flops=flops)
    return (out, lse) if return_lse else out


def patch_embed(images, descs_dev, cu_seqlens, total_tokens, patch_size, channels, weight, bias, pos_embedding,
                interpolate, out_dtype, patch_index=None):
    """K1 over a ragged image list.  `images` is only kept alive here; `descs_dev` is the device table."""
    _cuda(descs_dev, cu_seqlens, weight, pos_embedding, *images)
    E = weight.shape[0]
    out = torch.empty((total_tokens, E), dtype=out_dtype, device=weight.device)
    w, b, pe = _f32c(weight), _f32c(bias), _f32c(pos_embedding)
    lib = _lib.load()
    ws = None
    if out_dtype == torch.bfloat16 and not interpolate:     # tensor-core path: im2col + tcgen05 GEMM
        ws = torch.empty(lib.acai_patch_embed_workspace_bytes(total_tokens, patch_size, channels, E), dtype=torch.uint8,
                         device=weight.device)
    K = channels * patch_size * patch_size
    _call("acai_patch_embed_fwd", lib.acai_patch_embed_fwd, _p(descs_dev), len(images), _p(cu_seqlens), _p(patch_index), total_tokens,
                                        patch_size, channels, _p(w), _p(b), _p(pe), pe.shape[0], pe.shape[1],
                                        1 if interpolate else 0, _p(out), _DT[out_dtype], E, _p(ws),
                                        ws.numel() if ws is not None else 0, _stream(),
          flops=2.0 * total_tokens * K * E, nbytes=float(total_tokens) * (4 * K + 4 * E + out.element_size() * E))
    if ws is not None:
        _bump(2)  # im2col + weight cast + GEMM
    return out


class _PinnedRing:
    """Asynchronous upload of the small per-call host tables (cu_seqlens, image descriptors, padding masks).  A pageable
    `.to(device)` waits for everything already queued on the stream - at the top of a step that is the whole previous
    step - and the GPU then idles while the host catches up.  Here the table is copied into one of RING pinned buffers and
    uploaded with cudaMemcpyAsync; a buffer is only rewritten after the event recorded behind its last upload."""
    RING = 32

    def __init__(self):
        self.bufs, self.events, self.i = [None] * self.RING, [None] * self.RING, 0

    def upload(self, host: torch.Tensor, device) -> torch.Tensor:
        device = torch.device(device)
        if device.type != "cuda":
            return host.to(device)
        flat = host.contiguous().view(torch.uint8).reshape(-1)
        n = flat.numel()
        if n == 0:
            return torch.empty(host.shape, dtype=host.dtype, device=device)
        slot, self.i = self.i, (self.i + 1) % self.RING
        buf = self.bufs[slot]
        if self.events[slot] is not None:
            self.events[slot].synchronize()
        if buf is None or buf.numel() < n:
            buf = torch.empty(max(4096, 1 << (n - 1).bit_length()), dtype=torch.uint8).pin_memory()
            self.bufs[slot] = buf
        buf[:n].copy_(flat)
        with torch.cuda.device(device):
            dev = buf[:n].to(device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        self.events[slot] = ev
        return dev.view(host.dtype).view(host.shape)


_RING = _PinnedRing()


def upload_small(host: torch.Tensor, device) -> torch.Tensor:
    """Device copy of a small CPU tensor without stalling the host on the stream's queued work (see _PinnedRing)."""
    return _RING.upload(host, device)


def make_image_descs(images, patch_size, device):
    """Host-side table of acai_image_desc for a ragged list of (C,H,W) fp32 CUDA tensors -> device bytes."""
    arr = (ImageDesc * len(images))()
    for i, t in enumerate(images):
        H, W = t.shape[-2], t.shape[-1]
        arr[i] = ImageDesc(t.data_ptr(), H, W, H // patch_size, W // patch_size)
    host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
    return upload_small(host, device)


def decode_attn(q, k, v, kv_len, heads, head_dim, key_padding_mask=None, out_dtype=None, scale=None):
    """q:[B, H*d]; k,v: (B,H,T,d)-strided caches (acai_decode_attn)."""
    _cuda(q, k, v, key_padding_mask)
    B = q.shape[0]
    if k.stride(-1) != 1 or v.stride(-1) != 1 or k.stride() != v.stride() or k.dtype != v.dtype:
        raise AcaiError("decode_attn: k/v must share strides/dtype with a unit inner stride")
    if q.stride(-1) != 1:
        q = q.contiguous()
    mask = None
    if key_padding_mask is not None:
        mask = key_padding_mask.to(torch.uint8) if key_padding_mask.dtype != torch.uint8 else key_padding_mask
        mask = mask.contiguous()
    out_dtype = out_dtype or q.dtype
    out = torch.empty((B, heads * head_dim), dtype=out_dtype, device=q.device)
    lib = _lib.load()
    _call("acai_decode_attn", lib.acai_decode_attn, _p(q), _dt(q), q.stride(0), _p(k), _p(v), _dt(k), k.stride(0), k.stride(1),
                                    k.stride(2), B, heads, head_dim, int(kv_len), _p(mask),
                                    mask.stride(0) if mask is not None else 0,
                                    float(scale or 1.0 / math.sqrt(head_dim)), _p(out), _DT[out_dtype],
                                    out.stride(0), None, 0, _stream())
    return out


def kv_append(k_new, v_new, k_cache, v_cache, pos):
    """k_new/v_new: (B,H,s,d) strided views -> caches[:B,:,pos:pos+s] (acai_kv_append)."""
    _cuda(k_new, v_new, k_cache, v_cache)
    B, H, s, d = k_new.shape
    if k_new.stride(-1) != 1 or k_new.stride() != v_new.stride():
        k_new, v_new = k_new.contiguous(), v_new.contiguous()
    lib = _lib.load()
    _call("acai_kv_append", lib.acai_kv_append, _p(k_new), _p(v_new), _dt(k_new), k_new.stride(0), k_new.stride(1), k_new.stride(2),
                                  s, _p(k_cache), _p(v_cache), _dt(k_cache), k_cache.stride(0), k_cache.stride(1),
                                  k_cache.stride(2), B, H, d, int(pos), _stream())


def embed_tokens(ids, table, pos, out_dtype, pos_offset=0, period=0, pos_ids=None):
    """table[ids] + pos[...] (acai_embed_tokens); ids any integer dtype, any shape."""
    _cuda(ids, table, pos, pos_ids)
    flat = ids.reshape(-1)
    if flat.dtype != torch.int64:
        flat = flat.long()
    flat = flat.contiguous()
    dim = table.shape[1]
    out = torch.empty((flat.numel(), dim), dtype=out_dtype, device=table.device)
    t, p = _f32c(table), _f32c(pos)
    lib = _lib.load()
    _call("acai_embed_tokens", lib.acai_embed_tokens, _p(flat), flat.numel(), _p(t), _p(p), _p(pos_ids), int(pos_offset), int(period),
                                     _p(out), _DT[out_dtype], dim, _stream())
    return out.view(*ids.shape, dim)


def argmax_logprob(logits):
    """(next_token int64 [B], chosen log-prob fp32 [B]) from [B, V] logits (acai_argmax_logprob)."""
    _cuda(logits)
    l2, rows, vocab, ld = _rows2d(logits)
    nxt = torch.empty(rows, dtype=torch.int64, device=logits.device)
    lp = torch.empty(rows, dtype=torch.float32, device=logits.device)
    lib = _lib.load()
    _call("acai_argmax_logprob", lib.acai_argmax_logprob, _p(l2), _dt(l2), ld, rows, vocab, _p(nxt), _p(lp), _stream())
    return nxt, lp


def unpack_padded(packed, cu_seqlens, batch, max_len, fill=None, out_dtype=None):
    _cuda(packed, cu_seqlens, fill)
    dim = packed.shape[-1]
    packed = packed.contiguous()
    out_dtype = out_dtype or packed.dtype
    out = torch.empty((batch, max_len, dim), dtype=out_dtype, device=packed.device)
    f = _f32c(fill)
    lib = _lib.load()
    _call("acai_unpack_padded", lib.acai_unpack_padded, _p(packed), _dt(packed), _p(cu_seqlens), batch, max_len, dim, _p(f), _p(out),
                                      _DT[out_dtype], _stream())
    return out


def pack_padded(padded, cu_seqlens, total, out_dtype=None):
    _cuda(padded, cu_seqlens)
    B, L, dim = padded.shape
    padded = padded.contiguous()
    out_dtype = out_dtype or padded.dtype
    out = torch.empty((total, dim), dtype=out_dtype, device=padded.device)
    lib = _lib.load()
    _call("acai_pack_padded", lib.acai_pack_padded, _p(padded), _dt(padded), _p(cu_seqlens), B, L, dim, _p(out), _DT[out_dtype],
                                    _stream())
    return out


# ------------------------------------------------------------------------------------------ decode engine ops
def decode_workspace(batch, heads, max_kv_len, device):
    n = _lib.load().acai_decode_workspace_bytes(batch, heads, max_kv_len)
    return torch.zeros(n, dtype=torch.uint8, device=device)       # starts with the (self-resetting) arrival counters: must be zero


def decode_attn_paged(q, new_kv, kv_rows, heads, head_dim, max_kv_len, workspace, out_dtype, page_table=None,
                      seq_start=None, kv_lens=None, step=None, len_offset=0, scale=None):
    """Token-major (paged) decode attention (acai_decode_attn_paged).  q:[B, >=H*d] view, new_kv:[B, >=2*H*d] view or None,
    kv_rows:[rows, 2*H*d] cache."""
    _cuda(q, new_kv, kv_rows, page_table, seq_start, kv_lens, step, workspace)
    B = q.shape[0]
    out = torch.empty((B, heads * head_dim), dtype=out_dtype, device=q.device)
    lib = _lib.load()
    nbytes = float(kv_rows.element_size()) * 2 * heads * head_dim * B * max_kv_len  # upper bound; bench.py passes exact
    _call("acai_decode_attn_paged", lib.acai_decode_attn_paged, _p(q), q.stride(0), _p(new_kv),
          new_kv.stride(0) if new_kv is not None else 0, _p(kv_rows), _dt(kv_rows), _p(page_table),
          page_table.stride(0) if page_table is not None else 0, _p(seq_start), _p(kv_lens), _p(step), int(len_offset),
          int(max_kv_len), B, heads, head_dim, float(scale or 1.0 / math.sqrt(head_dim)), _p(out), _DT[out_dtype],
          out.stride(0), _p(workspace), workspace.numel(), kv_rows.shape[0], _stream(), nbytes=nbytes)
    if q.dtype != torch.bfloat16:
        _bump()  # fp32 caches: split-KV pass + a combine launch (the bf16 TMA kernel folds its partials itself)
    return out


def decode_embed(seqs, step, table, pos, out_dtype, out2_dtype=None):
    _cuda(seqs, step, table, pos)
    B, dim = seqs.shape[0], table.shape[1]
    out = torch.empty((B, dim), dtype=out_dtype, device=seqs.device)
    out2 = torch.empty((B, dim), dtype=out2_dtype, device=seqs.device) if out2_dtype is not None else None
    lib = _lib.load()
    _call("acai_decode_embed", lib.acai_decode_embed, _p(seqs), seqs.stride(0), _p(step), B, _p(_f32c(table)),
          _p(_f32c(pos)), _p(out), _DT[out_dtype], _p(out2), _DT[out2_dtype] if out2 is not None else 0, dim, _stream())
    return out if out2 is None else (out, out2)


def greedy_update(logits, step, seqs, log_probs, finished, eos_idx):
    _cuda(logits, step, seqs, log_probs, finished)
    l2, rows, vocab, ld = _rows2d(logits)
    lib = _lib.load()
    _call("acai_greedy_update", lib.acai_greedy_update, _p(l2), _dt(l2), ld, rows, vocab, _p(step), _p(seqs),
          seqs.stride(0), _p(log_probs), _p(finished), int(eos_idx), _stream())


def topk_sample_update(logits, step, seqs, log_probs, finished, eos_idx, top_k, temperature, seed=None, uniforms=None):
    """One GRPO rollout step on the device (acai_topk_sample_update): top-k filter, temperature softmax, one draw per
    row.  `seed`: 1-element int64 device tensor (hashed with row and step); `uniforms`: fp32 [rows, >= seqs.shape[1]]
    explicit draws in [0, 1) (tests)."""
    _cuda(logits, step, seqs, log_probs, finished, seed, uniforms)
    if seed is None and uniforms is None:
        raise AcaiError("topk_sample_update: pass a device seed tensor or explicit uniforms")
    l2, rows, vocab, ld = _rows2d(logits)
    if uniforms is not None and (uniforms.dtype != torch.float32 or uniforms.stride(-1) != 1 or uniforms.shape[0] != rows):
        raise AcaiError("topk_sample_update: uniforms must be fp32 [rows, steps] with a unit inner stride")
    if seed is not None and (seed.dtype != torch.int64 or seed.numel() != 1):
        raise AcaiError("topk_sample_update: seed must be a 1-element int64 device tensor")
    lib = _lib.load()
    _call("acai_topk_sample_update", lib.acai_topk_sample_update, _p(l2), _dt(l2), ld, rows, vocab, int(top_k), float(temperature),
          _p(uniforms), uniforms.stride(0) if uniforms is not None else 0, _p(seed), _p(step), _p(seqs), seqs.stride(0),
          _p(log_probs), _p(finished), int(eos_idx), _stream())


def advance_step(step):
    _cuda(step)
    lib = _lib.load()
    _call("acai_advance_step", lib.acai_advance_step, _p(step), _stream())


# ------------------------------------------------------------------------------------------ MAE glue
def patchify(images, descs_dev, cu_seqlens, total_tokens, patch_size, channels, out_dtype=torch.float32):
    """nn.Unfold patches of a ragged image list as packed rows [total, C*P*P] (acai_patchify)."""
    _cuda(descs_dev, cu_seqlens, *images)
    out = torch.empty((total_tokens, channels * patch_size * patch_size), dtype=out_dtype, device=cu_seqlens.device)
    _call("acai_patchify", _lib.load().acai_patchify, _p(descs_dev), len(images), _p(cu_seqlens), total_tokens, patch_size,
          channels, _p(out), _DT[out_dtype], _stream())
    return out


def mae_restore(latent, cu_keep, cu_full, ids_restore, patches_per_row, total_rows, mask_token, dec_pos_embedding, out_dtype):
    """MAE.prepare_for_decoder on packed rows (acai_mae_restore)."""
    _cuda(latent, cu_keep, cu_full, ids_restore, patches_per_row, mask_token, dec_pos_embedding)
    dim = latent.shape[-1]
    latent = latent.contiguous()
    out = torch.empty((total_rows, dim), dtype=out_dtype, device=latent.device)
    mt, pe = _f32c(mask_token).reshape(-1), _f32c(dec_pos_embedding)
    _call("acai_mae_restore", _lib.load().acai_mae_restore, _p(latent), _dt(latent), _p(cu_keep), _p(cu_full), _p(ids_restore),
          _p(patches_per_row), cu_full.numel() - 1, total_rows, _p(mt), _p(pe), pe.shape[1], _p(out), _DT[out_dtype], dim,
          _stream())
    return out


def mae_loss(pred, loss_mask, target):
    """(sum of per-patch losses over loss_mask, count) as a 2-element fp32 tensor, plus the per-row losses."""
    _cuda(pred, loss_mask, target)
    feats = pred.shape[-1]
    p2 = pred.reshape(-1, feats).contiguous()
    t2 = target.reshape(-1, feats).float().contiguous()
    m2 = loss_mask.reshape(-1).to(torch.uint8).contiguous()
    rows = p2.shape[0]
    row_loss = torch.empty(rows, dtype=torch.float32, device=pred.device)
    sums = torch.empty(2, dtype=torch.float32, device=pred.device)
    _call("acai_mae_loss_fwd", _lib.load().acai_mae_loss_fwd, _p(p2), _dt(p2), _p(t2), _p(m2), rows, feats, _p(row_loss),
          _p(sums), _stream())
    _bump()
    return sums, row_loss


# ------------------------------------------------------------------------------------------ training building blocks
def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def transpose_pad(x, pad_to=64):
    """[R, C] -> [C, ceil(R / pad_to) * pad_to] with zero padding (acai_transpose): the K-major operand of dW = dY^T X."""
    _cuda(x)
    if x.stride(1) != 1:
        x = x.contiguous()
    R, Cc = x.shape
    Rp = (R + pad_to - 1) // pad_to * pad_to
    out = torch.empty((Cc, Rp), dtype=x.dtype, device=x.device)
    _call("acai_transpose", _lib.load().acai_transpose, _p(x), _dt(x), x.stride(0), R, Cc, _p(out), Rp, _stream(),
          nbytes=2.0 * x.element_size() * R * Cc)
    return out


def wgrad_supported(dz, x, dw):
    return (dz.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and dw.dtype == torch.float32 and dz.dim() == 2 and
            x.dim() == 2 and dz.stride(1) == 1 and x.stride(1) == 1 and dw.stride(-1) == 1 and dz.stride(0) % 8 == 0 and
            x.stride(0) % 8 == 0 and dw.stride(0) % 4 == 0 and dz.data_ptr() % 16 == 0 and x.data_ptr() % 16 == 0 and
            dw.data_ptr() % 16 == 0 and x.shape[1] >= 8)


# bias gradients from the weight-gradient kernel's own pass over dz (acai_wgrad_bias_accumulate); ACAI_FUSE_BIAS_GRAD=0 runs
# acai_wgrad_accumulate + acai_colsum (the fixed-order two-stage column sum) instead
FUSE_BIAS_GRAD = os.environ.get("ACAI_FUSE_BIAS_GRAD", "1") != "0"


def wgrad_accumulate(dz, x, dw, dbias=None):
    """dw[n_out, k_in] += dz[T, n_out]^T @ x[T, k_in] in place on the tensor cores (acai_wgrad_accumulate), reading both
    operands token-major; falls back to explicit transposes + the forward GEMM when TMA cannot address the operands.
    `dbias` (fp32 [n_out], contiguous): += the column sums of dz, from the same pass over dz where the tensor-core kernel runs
    (acai_wgrad_bias_accumulate), by acai_colsum otherwise."""
    _cuda(dz, x, dw)
    T, n_out = dz.shape
    k_in = x.shape[1]
    if dw.shape != (n_out, k_in) or x.shape[0] != T:
        raise AcaiError("wgrad_accumulate: shape mismatch")
    if dbias is not None and (dbias.dtype != torch.float32 or dbias.shape != (n_out,) or not dbias.is_contiguous()):
        raise AcaiError("wgrad_accumulate: dbias must be a contiguous fp32 vector of n_out elements")
    if T == 0:
        return dw
    if wgrad_supported(dz, x, dw):
        if dbias is not None and FUSE_BIAS_GRAD:
            name = "acai_wgrad_bias_accumulate" if Profile.active is None else f"acai_wgrad_bias_accumulate[{n_out}x{k_in}x{T}]"
            _call(name, _lib.load().acai_wgrad_bias_accumulate, _p(dz), dz.stride(0), _p(x), x.stride(0), _dt(dz), _p(dw),
                  dw.stride(0), _p(dbias), n_out, k_in, T, _stream(), flops=2.0 * T * n_out * k_in)
            return dw
        name = "acai_wgrad_accumulate" if Profile.active is None else f"acai_wgrad_accumulate[{n_out}x{k_in}x{T}]"
        _call(name, _lib.load().acai_wgrad_accumulate, _p(dz), dz.stride(0), _p(x), x.stride(0), _dt(dz), _p(dw), dw.stride(0),
              n_out, k_in, T, _stream(), flops=2.0 * T * n_out * k_in)
    else:
        dw += linear(transpose_pad(dz), transpose_pad(x), out_dtype=torch.float32)
    if dbias is not None:
        colsum(dz, out=dbias, accumulate=True)
    return dw


# dX GEMMs read the forward's [out, in] weight copy as an MN-major B operand (acai_linear_dgrad); ACAI_DGRAD_KN=0 makes the
# callers transpose the weight every step and run acai_linear_fwd / acai_linear_gelu_bwd on it
DGRAD_KN = os.environ.get("ACAI_DGRAD_KN", "1") != "0"


def linear_dgrad(dz, weight, residual=None, z=None, dropout_p=0.0, dropout_seed=0):
    """dx = dz @ weight (+ residual | * gelu'(z) * dropout mask) with `weight` the forward's [out, in] bf16 copy
    (acai_linear_dgrad); None when the problem does not qualify (the caller runs the transposed-weight path)."""
    if not DGRAD_KN or dz.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16 or dz.dim() != 2 or weight.dim() != 2:
        return None
    _cuda(dz, weight, residual, z)
    M, K = dz.shape
    N = weight.shape[1]
    if weight.shape[0] != K or M == 0 or dz.stride(1) != 1 or weight.stride(1) != 1:
        return None
    for t in (residual, z):
        if t is not None and (t.dtype != torch.bfloat16 or t.shape != (M, N) or t.stride(1) != 1):
            return None
    if z is not None and (residual is not None or z.stride(0) != N):
        return None
    out = torch.empty((M, N), dtype=torch.bfloat16, device=dz.device)
    name = "acai_linear_dgrad" if Profile.active is None else \
        f"acai_linear_dgrad[{M}x{N}x{K}{',gelu_bwd' if z is not None else ''}{',res' if residual is not None else ''}]"
    ok = _try_call(name, _lib.load().acai_linear_dgrad, _p(dz), dz.stride(0), _p(weight), weight.stride(0), _p(residual),
                   residual.stride(0) if residual is not None else 0, _p(z), z.stride(0) if z is not None else 0, _p(out),
                   out.stride(0), M, N, K, float(dropout_p), C.c_uint64(dropout_seed), _stream(), flops=2.0 * M * N * K)
    return out if ok else None


ERR_UNSUPPORTED = -2
# GELU (+ dropout) forward / backward inside the tcgen05 GEMM epilogues (acai_linear_gelu_train / acai_linear_gelu_bwd).
# With the A&S erf form (~27 extra instructions per element on eight epilogue warps that are already the critical path at
# K = 768 / 1024) this was slower than the standalone 16-byte GELU kernels (208.7 vs 203.4 ms, profiles/r1g_gelu_epilogue.md);
# with the tanh-fit forms of round 1i (gelu_fast / gelu_grad_fast, 10 / 14 instructions) it wins: teacher-forced step
# 178.3 -> 174.6 ms, MAE step unchanged.  ACAI_FUSE_GELU_GEMM=0 selects the standalone kernels.
FUSE_GELU_GEMM = os.environ.get("ACAI_FUSE_GELU_GEMM", "1") != "0"


def _try_call(name, fn, *args, flops=0.0, nbytes=0.0):
    """Like _call, but returns False (nothing launched) when the entry point reports ACAI_ERR_UNSUPPORTED."""
    prof = Profile.active
    if prof is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    rc = fn(*args)
    if rc == ERR_UNSUPPORTED:
        return False
    _lib.check(rc, name)
    if prof is not None:
        e1.record()
        prof.records.append((name, e0, e1, flops, nbytes))
    _bump()
    return True


def _gemm_fusable(*ts):
    return all(t.dtype == torch.bfloat16 and t.dim() == 2 and t.stride(1) == 1 and t.stride(0) % 8 == 0 and
                                  t.data_ptr() % 16 == 0 for t in ts)


def splitk_workspace(M, N, K, device):
    """Workspace large enough for linear_splitk on [M, K] x [N, K] (acai_linear_splitk_workspace_bytes)."""
    return _ws(_lib.load().acai_linear_splitk_workspace_bytes(int(M), int(N), int(K)), device)


def linear_splitk(x, weight, bias=None, act=ACT_NONE, residual=None, ln=None, out_dtype=None, workspace=None,
                  out2_dtype=None):
    """LayerNorm?(act(x @ weight.T + bias) + residual) for a SMALL number of rows (decode step): split-K tcgen05 GEMM +
    one reduce kernel (acai_linear_splitk).  `ln` = (gamma, beta, eps) or None.  Returns None when the problem does not
    qualify (the caller runs linear [+ layernorm]).  `workspace` (uint8) is reused when large enough.  With `out2_dtype`
    the result is (out, out2): the same rows in a second dtype (fp32 residual stream + bf16 GEMM operand)."""
    _cuda(x, weight, bias, residual)
    if x.dim() != 2 or not _gemm_fusable(x, weight):
        return None
    M, K = x.shape
    N = weight.shape[0]
    if M == 0 or M > 256:
        return None
    lib = _lib.load()
    need = lib.acai_linear_splitk_workspace_bytes(M, N, K)
    if workspace is None or workspace.numel() < need:
        workspace = _ws(need, x.device)
    out_dtype = out_dtype or x.dtype
    out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    out2 = torch.empty((M, N), dtype=out2_dtype, device=x.device) if out2_dtype is not None else None
    r2, ldr, rdt = None, 0, 0
    if residual is not None:
        r2, rm, rn, ldr = _rows2d(residual)
        if rm != M or rn != N:
            raise AcaiError("linear_splitk: residual shape mismatch")
        rdt = _dt(r2)
    g, b2, eps = (None, None, 0.0) if ln is None else (_f32c(ln[0]), _f32c(ln[1]), float(ln[2]))
    name = "acai_linear_splitk" if Profile.active is None else \
        f"acai_linear_splitk[{M}x{N}x{K}{',gelu' if act else ''}{',res' if r2 is not None else ''}{',ln' if ln else ''}]"
    ok = _try_call(name, lib.acai_linear_splitk, _p(x), x.stride(0), _p(weight), weight.stride(0), _p(_f32c(bias)), _p(r2), rdt, ldr,
                   _p(g), _p(b2), eps, _p(out), _DT[out_dtype], out.stride(0), _p(out2),
                   _DT[out2_dtype] if out2 is not None else 0, N, M, N, K, act, _p(workspace), workspace.numel(),
                   _stream(), flops=2.0 * M * N * K, nbytes=2.0 * (M * K + N * K + M * N))
    if ok:
        _bump()                   # two kernels: the split-K GEMM and the reduce
    if not ok:
        return None
    return out if out2 is None else (out, out2)


FUSE_LINEAR_LN = os.environ.get("ACAI_FUSE_LINEAR_LN", "1") != "0"


def linear_layernorm(x, weight, bias, residual, gamma, beta, eps, want_f32):
    """LayerNorm(x @ weight.T + bias + residual) * gamma + beta as one cluster kernel (acai_linear_layernorm): returns
    (out_bf16, out_f32 or None), or None when the problem does not qualify (the caller runs linear + layernorm)."""
    if not FUSE_LINEAR_LN:
        return None
    _cuda(x, weight, bias, residual, gamma, beta)
    if x.dim() != 2 or residual.dim() != 2 or not _gemm_fusable(x, weight) or residual.stride(1) != 1:
        return None
    M, K = x.shape
    N = weight.shape[0]
    if M == 0 or N not in (512, 768, 1024) or residual.shape != (M, N):
        return None
    out16 = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    out32 = torch.empty((M, N), dtype=torch.float32, device=x.device) if want_f32 else None
    lib = _lib.load()
    name = "acai_linear_layernorm" if Profile.active is None else f"acai_linear_layernorm[{M}x{N}x{K}{',f32' if want_f32 else ''}]"
    ok = _try_call(name, lib.acai_linear_layernorm, _p(x), x.stride(0), _p(weight), weight.stride(0), _p(_f32c(bias)), _p(residual),
                   _dt(residual), residual.stride(0), _p(_f32c(gamma)), _p(_f32c(beta)), float(eps), _p(out16), N, _p(out32), N,
                   M, N, K, _stream(), flops=2.0 * M * N * K,
                   nbytes=2.0 * (M * K + N * K) + M * N * (residual.element_size() + 2 + (4 if want_f32 else 0)))
    return (out16, out32) if ok else None


def linear_gelu_train(x, weight, bias, dropout_p=0.0, dropout_seed=0):
    """(z, h) with z = x @ weight.T + bias and h = dropout(gelu(z)), both bf16, from one tcgen05 GEMM
    (acai_linear_gelu_train); None when the operands do not qualify (the caller runs linear + gelu)."""
    _cuda(x, weight, bias)
    M, K = x.shape
    N = weight.shape[0]
    if not _gemm_fusable(x, weight) or M == 0:
        return None
    z = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    h = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    b = _f32c(bias)
    name = "acai_linear_gelu_train" if Profile.active is None else f"acai_linear_gelu_train[{M}x{N}x{K}]"
    ok = _try_call(name, _lib.load().acai_linear_gelu_train, _p(x), x.stride(0), _p(weight), weight.stride(0), _p(b), _p(z), N,
                   _p(h), N, M, N, K, float(dropout_p), int(dropout_seed), _stream(), flops=2.0 * M * N * K,
                   nbytes=2.0 * (M * K + N * K + 2 * M * N))
    return (z, h) if ok else None


FUSE_DROPOUT_GEMM = os.environ.get("ACAI_FUSE_DROPOUT_GEMM", "1") != "0"


def linear_dropout(x, weight, bias, residual, dropout_p, dropout_seed):
    """residual + dropout(x @ weight.T + bias), bf16, from one tcgen05 GEMM (acai_linear_dropout; the mask is `dropout`'s on
    the dense [M, N] output); None when the operands do not qualify (the caller runs linear + dropout)."""
    _cuda(x, weight, bias, residual)
    M, K = x.shape
    N = weight.shape[0]
    if not FUSE_DROPOUT_GEMM or not _gemm_fusable(x, weight) or M == 0:
        return None
    if residual is not None and (residual.dim() != 2 or residual.shape != (M, N) or residual.stride(1) != 1 or
                                 residual.dtype not in (torch.bfloat16, torch.float32)):
        return None
    out = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    b = _f32c(bias)
    name = "acai_linear_dropout" if Profile.active is None else f"acai_linear_dropout[{M}x{N}x{K}]"
    ok = _try_call(name, _lib.load().acai_linear_dropout, _p(x), x.stride(0), _p(weight), weight.stride(0), _p(b), _p(residual),
                   _dt(residual) if residual is not None else 1, residual.stride(0) if residual is not None else 0, _p(out), N,
                   M, N, K, float(dropout_p), C.c_uint64(dropout_seed), _stream(), flops=2.0 * M * N * K,
                   nbytes=2.0 * (M * K + N * K + 2 * M * N))
    return out if ok else None


def linear_gelu_bwd(dy, weight_t, z, dropout_p=0.0, dropout_seed=0):
    """dz = (dy @ weight_t.T) * gelu'(z) * dropout_mask from one tcgen05 GEMM (acai_linear_gelu_bwd); None when the
    operands do not qualify (the caller runs linear + gelu_bwd)."""
    _cuda(dy, weight_t, z)
    M, K = dy.shape
    N = weight_t.shape[0]
    if not _gemm_fusable(dy, weight_t, z) or M == 0 or z.shape != (M, N) or z.stride(0) != N:
        return None
    dz = torch.empty((M, N), dtype=torch.bfloat16, device=dy.device)
    name = "acai_linear_gelu_bwd" if Profile.active is None else f"acai_linear_gelu_bwd[{M}x{N}x{K}]"
    ok = _try_call(name, _lib.load().acai_linear_gelu_bwd, _p(dy), dy.stride(0), _p(weight_t), weight_t.stride(0), _p(z), N, _p(dz),
                   N, M, N, K, float(dropout_p), int(dropout_seed), _stream(), flops=2.0 * M * N * K,
                   nbytes=2.0 * (M * K + N * K + 2 * M * N))
    return dz if ok else None


def colsum(x, out=None, accumulate=False):
    """fp32 column sums of a 2-D tensor in a fixed order (acai_colsum)."""
    _cuda(x, out)
    if x.stride(1) != 1:
        x = x.contiguous()
    R, Cc = x.shape
    if out is None:
        out = torch.empty(Cc, dtype=torch.float32, device=x.device)
    lib = _lib.load()
    ws = _ws(lib.acai_colsum_workspace_bytes(R, Cc), x.device)
    _call("acai_colsum", lib.acai_colsum, _p(x), _dt(x), x.stride(0), R, Cc, _p(out), 1 if accumulate else 0, _p(ws),
          ws.numel(), _stream())
    _bump()
    return out


def gelu(z, out_dtype=None, dropout_p=0.0, dropout_seed=0):
    """dropout(gelu(z)) - exact-erf GELU, optional fused nn.Dropout (acai_gelu_fwd)."""
    _cuda(z)
    z = z.contiguous()
    out = torch.empty(z.shape, dtype=out_dtype or z.dtype, device=z.device)
    _call("acai_gelu_fwd", _lib.load().acai_gelu_fwd, _p(z), _dt(z), _p(out), _dt(out), z.numel(), float(dropout_p),
          C.c_uint64(dropout_seed), _stream())
    return out


def gelu_bwd(dy, z, out_dtype=None, dropout_p=0.0, dropout_seed=0):
    _cuda(dy, z)
    dy, z = dy.contiguous(), z.contiguous()
    dz = torch.empty(z.shape, dtype=out_dtype or z.dtype, device=z.device)
    _call("acai_gelu_bwd", _lib.load().acai_gelu_bwd, _p(dy), _dt(dy), _p(z), _dt(z), _p(dz), _dt(dz), z.numel(), float(dropout_p),
          C.c_uint64(dropout_seed), _stream())
    return dz


def layernorm_bwd(dy, x, residual, gamma, eps, dv_dtype=None, dgamma_acc=None, dbeta_acc=None):
    """-> (dv [rows, cols], dgamma fp32, dbeta fp32) for y = LN(x + residual) * gamma + beta.  With `dgamma_acc` /
    `dbeta_acc` (fp32 [cols], both or neither) the parameter gradients are ADDED into them instead (no temporaries, no
    add kernels) and returned as None."""
    _cuda(dy, x, residual, gamma)
    cols = x.shape[-1]
    x2, dy2 = x.reshape(-1, cols).contiguous(), dy.reshape(-1, cols).contiguous()
    r2 = residual.reshape(-1, cols).contiguous() if residual is not None else None
    rows = x2.shape[0]
    dv = torch.empty((rows, cols), dtype=dv_dtype or x.dtype, device=x.device)
    acc = dgamma_acc is not None and dbeta_acc is not None
    if acc:
        if not (dgamma_acc.dtype == dbeta_acc.dtype == torch.float32 and dgamma_acc.is_contiguous() and dbeta_acc.is_contiguous()
                and dgamma_acc.numel() == cols and dbeta_acc.numel() == cols):
            raise AcaiError("layernorm_bwd: accumulation targets must be contiguous fp32 [cols]")
        dg, db = dgamma_acc, dbeta_acc
    else:
        dg = torch.empty(cols, dtype=torch.float32, device=x.device)
        db = torch.empty(cols, dtype=torch.float32, device=x.device)
    lib = _lib.load()
    ws = _ws(lib.acai_layernorm_bwd_workspace_bytes(rows, cols), x.device)
    _call("acai_layernorm_bwd", lib.acai_layernorm_bwd, _p(dy2), _dt(dy2), _p(x2), _dt(x2), _p(r2),
          _dt(r2) if r2 is not None else 0, _p(_f32c(gamma)), float(eps), _p(dv), _dt(dv), _p(dg), _p(db), 1 if acc else 0, rows,
          cols, _p(ws), ws.numel(), _stream())
    _bump(1)
    return (dv.view(x.shape), None, None) if acc else (dv.view(x.shape), dg, db)


def attn_varlen_bwd(q, k, v, out, dout, lse, cu_q, cu_k, max_q, max_k, heads, head_dim, causal, dq, dk, dv, scale=None,
                    dropout_p=0.0, dropout_seed=0):
    """Writes dq/dk/dv (row-strided 2-D views, e.g. slices of a fused d(QKV) buffer) - acai_attn_varlen_bwd."""
    _cuda(q, k, v, out, dout, lse, cu_q, cu_k, dq, dk, dv)
    dout = dout if dout.stride(1) == 1 else dout.contiguous()
    if dout.dtype != q.dtype:
        dout = dout.to(q.dtype)
    delta = torch.empty(heads * q.shape[0], dtype=torch.float32, device=q.device)
    _call("acai_attn_varlen_bwd", _lib.load().acai_attn_varlen_bwd, _p(q), _p(k), _p(v), _dt(q), q.stride(0), k.stride(0),
          v.stride(0), _p(out), out.stride(0), _p(dout), dout.stride(0), _p(lse), _p(cu_q), _p(cu_k), cu_q.numel() - 1,
          int(max_q), int(max_k), q.shape[0], k.shape[0], heads, head_dim, 1 if causal else 0, float(scale or 1.0 / math.sqrt(head_dim)),
          _p(dq), _p(dk), _p(dv), _dt(dq), dq.stride(0), dk.stride(0), dv.stride(0), _p(delta), float(dropout_p),
          C.c_uint64(dropout_seed), _stream())
    _bump(2)


def ce_loss(logits, targets, ignore_index, dlogits_scale=None, dlogits_dtype=None, label_smoothing=0.0):
    """OMRCELoss pieces (acai_ce_loss).  Forward (dlogits_scale None): returns sums[2] = {sum, count}.
    Backward (dlogits_scale = 1-element fp32 device tensor): returns dlogits = scale * (softmax - q), q the smoothed
    target distribution."""
    _cuda(logits, targets, dlogits_scale)
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)
    if l2.stride(1) != 1:
        l2 = l2.contiguous()
    t = targets.reshape(-1).to(torch.int64).contiguous()
    rows = l2.shape[0]
    lib = _lib.load()
    if dlogits_scale is None:
        row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        valid = torch.empty(rows, dtype=torch.uint8, device=logits.device)
        sums = torch.empty(2, dtype=torch.float32, device=logits.device)
        _call("acai_ce_loss", lib.acai_ce_loss, _p(l2), _dt(l2), l2.stride(0), _p(t), rows, V, int(ignore_index), _p(row_loss),
              _p(valid), _p(sums), None, 0, 0, None, float(label_smoothing), _stream())
        _bump()
        return sums
    dl = torch.empty((rows, V), dtype=dlogits_dtype or logits.dtype, device=logits.device)
    _call("acai_ce_loss", lib.acai_ce_loss, _p(l2), _dt(l2), l2.stride(0), _p(t), rows, V, int(ignore_index), None, None, None,
          _p(dl), _dt(dl), V, _p(dlogits_scale), float(label_smoothing), _stream())
    return dl.view(logits.shape)


def scatter_add_rows(src, index, dst, scale=1.0):
    """dst[index[r]] += scale * src[r] (fp32 dst, rows with index < 0 skipped) - acai_scatter_add_rows."""
    _cuda(src, index, dst)
    dim = src.shape[-1]
    s2 = src.reshape(-1, dim)
    if s2.stride(1) != 1:
        s2 = s2.contiguous()
    idx = index.reshape(-1).to(torch.int64).contiguous()
    _call("acai_scatter_add_rows", _lib.load().acai_scatter_add_rows, _p(s2), _dt(s2), s2.stride(0), _p(idx), s2.shape[0], dim,
          _p(dst), float(scale), _stream())
    return dst


def gather_rows(src, index, out_dtype=None):
    """out[r] = src[index[r]] (0 where index < 0) - acai_gather_rows."""
    _cuda(src, index)
    dim = src.shape[-1]
    s2 = src.reshape(-1, dim).contiguous()
    idx = index.reshape(-1).to(torch.int64).contiguous()
    out = torch.empty((idx.numel(), dim), dtype=out_dtype or src.dtype, device=src.device)
    _call("acai_gather_rows", _lib.load().acai_gather_rows, _p(s2), _dt(s2), _p(idx), idx.numel(), dim, _p(out), _dt(out), _stream())
    return out


def mae_loss_bwd(pred, loss_mask, target, scale, dpred_dtype=None):
    _cuda(pred, loss_mask, target, scale)
    feats = pred.shape[-1]
    p2 = pred.reshape(-1, feats).contiguous()
    t2 = target.reshape(-1, feats).float().contiguous()
    m2 = loss_mask.reshape(-1).to(torch.uint8).contiguous()
    dp = torch.empty(p2.shape, dtype=dpred_dtype or pred.dtype, device=pred.device)
    _call("acai_mae_loss_bwd", _lib.load().acai_mae_loss_bwd, _p(p2), _dt(p2), _p(t2), _p(m2), p2.shape[0], feats, _p(scale),
          _p(dp), _dt(dp), _stream())
    return dp.view(pred.shape)


# ------------------------------------------------------------------------------------------ scheduled sampling
def gumbel_softmax(logits, noise, tau, hard, out_dtype=None):
    """-> (y, y_soft fp32) over the last dim; `noise` is the Gumbel(0,1) draw of the same shape (acai_gumbel_softmax_fwd)."""
    _cuda(logits, noise)
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)
    if l2.stride(1) != 1:
        l2 = l2.contiguous()
    n2 = noise.reshape(-1, V).float().contiguous()
    y_soft = torch.empty(l2.shape, dtype=torch.float32, device=logits.device)
    y = torch.empty(l2.shape, dtype=out_dtype or logits.dtype, device=logits.device)
    _call("acai_gumbel_softmax_fwd", _lib.load().acai_gumbel_softmax_fwd, _p(l2), _dt(l2), l2.stride(0), _p(n2), l2.shape[0], V,
          float(tau), 1 if hard else 0, _p(y_soft), _p(y), _dt(y), _stream())
    return y.view(logits.shape), y_soft.view(logits.shape)


def gumbel_softmax_bwd(dy, y_soft, tau, out_dtype):
    _cuda(dy, y_soft)
    V = y_soft.shape[-1]
    d2 = dy.reshape(-1, V).contiguous()
    dl = torch.empty(d2.shape, dtype=out_dtype, device=dy.device)
    _call("acai_gumbel_softmax_bwd", _lib.load().acai_gumbel_softmax_bwd, _p(d2), _dt(d2), _p(y_soft), d2.shape[0], V, float(tau),
          _p(dl), _dt(dl), _stream())
    return dl.view(y_soft.shape)


def mix_rows(b, a=None, sel=None, pos=None, period=0, out_dtype=None):
    """out[r] = (sel[r] >= 0 ? a[sel[r]] : b[r]) + pos[r % period] over rows of the last dim (acai_mix_rows)."""
    _cuda(a, sel, b, pos)
    dim = b.shape[-1]
    b2 = b.reshape(-1, dim).contiguous()
    a2 = a.reshape(-1, dim).contiguous() if a is not None else None
    s = sel.reshape(-1).to(torch.int64).contiguous() if sel is not None else None
    out = torch.empty(b2.shape, dtype=out_dtype or b.dtype, device=b.device)
    _call("acai_mix_rows", _lib.load().acai_mix_rows, _p(a2), _dt(a2) if a2 is not None else 0, _p(s), _p(b2), _dt(b2),
          _p(_f32c(pos)), int(period), b2.shape[0], dim, _p(out), _dt(out), _stream())
    return out.view(b.shape)


# ------------------------------------------------------------------------------------------ dropout
class DropoutState:
    """Seeds for the counter-based dropout masks: one 64-bit seed per dropout call, derived on the host from a base
    seed and a call counter (no device synchronisation).  `manual_seed` makes a run reproducible."""
    base = 0x5EED_ACA1_0000_0001
    counter = 0

    @classmethod
    def manual_seed(cls, seed: int):
        cls.base, cls.counter = int(seed) & 0xFFFFFFFFFFFFFFFF, 0

    @classmethod
    def next_seed(cls) -> int:
        cls.counter += 1
        x = (cls.base + 0x9E3779B97F4A7C15 * cls.counter) & 0xFFFFFFFFFFFFFFFF      # splitmix64 step
        x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return x ^ (x >> 31)


def dropout(x, p, seed, residual=None, out_dtype=None):
    """residual + keep(seed) * x / (1 - p)  (acai_dropout); the same call on a gradient is the backward."""
    _cuda(x, residual)
    x = x.contiguous()
    r = residual.contiguous() if residual is not None else None
    out = torch.empty(x.shape, dtype=out_dtype or x.dtype, device=x.device)
    _call("acai_dropout", _lib.load().acai_dropout, _p(x), _dt(x), _p(r), _dt(r) if r is not None else 0, x.numel(), float(p),
          C.c_uint64(seed), _p(out), _dt(out), _stream())
    return out


def dropout_mask(n, p, seed, device):
    keep = torch.empty(n, dtype=torch.uint8, device=device)
    _call("acai_dropout_mask", _lib.load().acai_dropout_mask, n, float(p), C.c_uint64(seed), _p(keep), _stream())
    return keep.bool()


def attn_dropout_mask(p, seed, heads, total_q, max_k, device):
    """keep[h, q, k] (bool) of the attention-probability dropout for (p, seed) - what the attention kernels regenerate."""
    keep = torch.empty((heads, total_q, max_k), dtype=torch.uint8, device=device)
    _call("acai_attn_dropout_mask", _lib.load().acai_attn_dropout_mask, float(p), C.c_uint64(seed), heads, total_q, max_k,
          _p(keep), _stream())
    return keep.bool()
