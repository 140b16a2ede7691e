"""Drop-in for `acai_omr/models/kv_caching.py` (jsnchon/acai-omr @ 5819ae3) on sm_100a kernels.

Same classes, constructor arguments, method signatures, buffer names and `state_dict` keys as the
reference (kv_caching.py:5-302); the arithmetic runs in libacai_b200.so:
  KVCache.update                          -> acai_kv_append            (kv_caching.py:57-109)
  CachedMultiheadAttention.cached_forward -> acai_decode_attn + acai_linear_fwd (:123-140)
  CachedTransformerDecoderLayer.cached_forward -> the above + acai_layernorm_fwd (:190-223)
  MemoryCache.cache_memory_keys_and_vals  -> acai_linear_fwd           (:235-253)
Differences that are deliberate: no host sync per update (the fill position is tracked on the host,
`size` costs nothing), and `reset()` does not memset 2 x (B,H,T,d) per layer: stale rows are never
read because every read is bounded by the fill position.
"""
from __future__ import annotations

import torch
from torch import nn

from .. import ops
from ..layers import (LayerNorm, Linear, MultiheadAttention, TransformerDecoder, TransformerDecoderLayer,
                      compute_dtype, norm_pair, stream_dtype, w_)


class KVCache(nn.Module):
    """Pre-allocated (B, H, T, d) key/value buffers (kv_caching.py:26-45)."""

    def __init__(self, max_batch_size: int, max_seq_len: int, num_kv_heads: int, head_dim: int, dtype: torch.dtype) -> None:
        super().__init__()
        cache_shape = (max_batch_size, num_kv_heads, max_seq_len, head_dim)
        self.register_buffer("k_cache", torch.zeros(cache_shape, dtype=dtype), persistent=False)
        self.register_buffer("v_cache", torch.zeros(cache_shape, dtype=dtype), persistent=False)
        self.max_batch_size = max_batch_size
        self.max_seq_len = max_seq_len
        self._pos = 0

    @property
    def cache_pos(self) -> torch.Tensor:
        """The reference keeps `arange(T) + size` as a buffer (kv_caching.py:42-44); derived here."""
        return torch.arange(self._pos, self._pos + self.max_seq_len, device=self.k_cache.device)

    def reset(self) -> None:
        self._pos = 0

    @property
    def size(self) -> int:
        return self._pos

    def update(self, k_val: torch.Tensor, v_val: torch.Tensor):
        cur_bsz, _, seq_len, _ = k_val.shape
        if cur_bsz > self.k_cache.shape[0]:
            raise ValueError(
                f"The current cache has been setup with a max batch size of {self.k_cache.shape[0]}"
                f", but found new key tensors with batch size {k_val.shape[0]}!")
        assert (self._pos + seq_len) <= self.k_cache.shape[2]
        ops.kv_append(k_val, v_val, self.k_cache, self.v_cache, self._pos)
        self._pos += seq_len
        return self.k_cache[:cur_bsz, :, :self._pos], self.v_cache[:cur_bsz, :, :self._pos]


class CachedMultiheadAttention(MultiheadAttention):
    def cached_forward(self, q_t, K_t, V_t, memory_key_padding_mask=None, residual=None):
        """q_t (B,H,1,d); K_t/V_t (B,H,T,d) -> out_proj(attention) (B,1,D)  (kv_caching.py:123-140).
        `residual` (B,1,D), if given, is added in the projection's epilogue."""
        B, H, _, d = q_t.shape
        q2 = q_t.reshape(B, H * d)
        a = ops.decode_attn(q2, K_t, V_t, K_t.shape[2], H, d, key_padding_mask=memory_key_padding_mask,
                            out_dtype=q_t.dtype)
        dt = a.dtype
        out = ops.linear(a, w_(self.out_proj.weight, dt), self.out_proj.bias,
                         residual=None if residual is None else residual.reshape(B, H * d),
                         out_dtype=dt if residual is None else residual.dtype)      # the sum joins the residual stream
        return out.view(B, 1, H * d)


class CachedTransformerDecoderLayer(TransformerDecoderLayer):
    def __init__(self, d_model: int, nhead: int, dim_feedforward: int = 2048, dropout: float = 0.1,
                 activation="gelu", layer_norm_eps: float = 1e-5, batch_first: bool = True,
                 norm_first: bool = False, bias: bool = True, device=None, dtype=None):
        super().__init__(d_model, nhead, dim_feedforward, dropout, activation, layer_norm_eps, batch_first,
                         norm_first, bias, device, dtype)
        self.hidden_dim = d_model
        self.num_heads = nhead
        self.head_dim = self.self_attn.head_dim

    def _make_attention(self, d_model, nhead, dropout, **kw):
        return CachedMultiheadAttention(d_model, nhead, dropout=dropout, **kw)

    def cached_forward(self, tgt_t, self_attn_kv_cache: KVCache, cached_kv_mem, memory_key_padding_mask=None):
        """One decode step for one layer (kv_caching.py:190-223); tgt_t (B,1,D)."""
        dt = compute_dtype()
        sdt = stream_dtype(dt)               # the residual stream (fp32 under bf16 autocast, as in the reference); xb = its
        x = tgt_t if tgt_t.dtype == sdt else tgt_t.to(sdt)          # copy in the GEMM operand dtype
        xb = x if sdt == dt else x.to(dt)
        B, D, H, d = tgt_t.shape[0], self.hidden_dim, self.num_heads, self.head_dim
        qkv = ops.linear(xb, w_(self.self_attn.in_proj_weight, dt), self.self_attn.in_proj_bias, out_dtype=dt)
        q_t, k_t, v_t = (c.view(B, H, 1, d) for c in qkv.chunk(3, dim=-1))
        K_t, V_t = self_attn_kv_cache.update(k_t, v_t)
        y = self.self_attn.cached_forward(q_t, K_t, V_t, residual=x)
        x, xb = norm_pair(self.norm1, y, dt)

        q_c = ops.linear(xb, w_(self.multihead_attn.in_proj_weight, dt, rows=(0, D)),
                         self.multihead_attn.in_proj_bias[:D], out_dtype=dt).view(B, H, 1, d)
        K_cross, V_cross = cached_kv_mem
        y = self.multihead_attn.cached_forward(q_c, K_cross, V_cross, memory_key_padding_mask=memory_key_padding_mask,
                                               residual=x)
        x, xb = norm_pair(self.norm2, y, dt)
        x, xb = self.mlp_packed(x.view(B, D), xb.view(B, D))
        return x.view(B, 1, D)


class MemoryCache(nn.Module):
    """Per-layer cross-attention K/V of the encoder memory (kv_caching.py:227-256)."""

    def __init__(self):
        super().__init__()
        self.register_buffer("K_cross", None, persistent=False)
        self.register_buffer("b_kv", None, persistent=False)
        self.V_cross = None

    def cache_memory_keys_and_vals(self, memory: torch.Tensor, layer: CachedTransformerDecoderLayer, cache_dtype=None):
        B, S, D = memory.shape
        H, d = layer.num_heads, layer.head_dim
        dt = compute_dtype()
        if memory.dtype != dt:
            memory = memory.to(dt)
        kv = ops.linear(memory, w_(layer.multihead_attn.in_proj_weight, dt, rows=(D, 3 * D)),
                        layer.multihead_attn.in_proj_bias[D:], out_dtype=cache_dtype or dt)
        K, V = kv.chunk(2, dim=-1)
        self.K_cross = K.view(B, S, H, d).transpose(1, 2)
        self.V_cross = V.view(B, S, H, d).transpose(1, 2)

    def get_cached_keys_and_vals(self):
        return self.K_cross, self.V_cross


class CachedTransformerDecoder(TransformerDecoder):
    def __init__(self, decoder_layer: CachedTransformerDecoderLayer, num_layers: int, max_batch_size: int,
                 max_decoder_seq_len: int, cache_dtype, norm: nn.Module = None):
        assert isinstance(decoder_layer, CachedTransformerDecoderLayer), \
            "Can't use uncached TransformerDecoderLayer in a cached TransformerDecoder"
        proto = decoder_layer

        def clone():
            layer = CachedTransformerDecoderLayer(
                proto.hidden_dim, proto.num_heads, proto.linear1.out_features, proto.dropout_p,
                layer_norm_eps=proto.norm1.eps, device=proto.linear1.weight.device, dtype=proto.linear1.weight.dtype)
            layer.load_state_dict(proto.state_dict())
            return layer

        super().__init__(clone, num_layers, norm)
        self.self_attn_caches = nn.ModuleList([
            KVCache(max_batch_size, max_decoder_seq_len, layer.num_heads, layer.head_dim, dtype=cache_dtype)
            for layer in self.layers])
        self.cross_attn_caches = nn.ModuleList([MemoryCache() for _ in self.layers])

    def prepare_caches(self, encoder_memory):
        for c in self.self_attn_caches:
            c.reset()
        for i, c in enumerate(self.cross_attn_caches):
            c.cache_memory_keys_and_vals(encoder_memory, self.layers[i])

    def cached_generate(self, embedding_t, memory_key_padding_mask=None):
        x = embedding_t
        for i, layer in enumerate(self.layers):
            x = layer.cached_forward(x, self.self_attn_caches[i], self.cross_attn_caches[i].get_cached_keys_and_vals(),
                                     memory_key_padding_mask=memory_key_padding_mask)
        if self.norm is not None:
            x = self.norm(x, out_dtype=compute_dtype())
        elif x.dtype != compute_dtype():
            x = x.to(compute_dtype())
        return x
