"""KV-cached greedy decode engine: ViTOMR.cached_greedy_generate's loop (models.py:562-615) with every step
captured ONCE as a CUDA graph and replayed, no host sync per step.

Layout in HBM (dtype = compute dtype, bf16 under autocast / fp32 otherwise):
  self cache   pool[layer]  : [B * pages_per_seq * 64 rows, 2*D]  token-major rows [K | V], 64-token pages, page table [B, pages]
  memory cache cross[layer] : [sum S_b rows, 2*D]  = the cross-KV projection GEMM's output, packed (no padding rows)
  per step     x [B, D], qkv [B, 3D], mlp hidden [B, M], logits [B, V]
Per step and layer: qkv GEMM -> paged split-KV attention (appends this step's K|V row) -> out-proj GEMM (+residual)
-> LN -> cross-q GEMM -> attention over the memory cache -> out-proj GEMM (+residual) -> LN -> MLP GEMMs -> LN.
Precision of the bf16 mode = what torch.autocast does to the reference's cached loop (vitomr_inference.py:82-84): Linear
and attention operands / results in bf16, but the RESIDUAL STREAM in fp32 - the embedding rows are fp32 parameters, a
bf16 sub-layer output added to them promotes to fp32, and LayerNorm runs (and returns) fp32 under autocast.  So the step
carries x twice: `x` fp32 [B, D] (residual operand of the next reduce / LayerNorm) and `xb` bf16 (operand of the next
GEMM), both written by the same kernel.  ACAI_DECODE_STREAM=bf16 restores round 1's single bf16 stream (1.7x the logit
error against the oracle, profiles/r2a_parity.md).
The step index lives in device memory; the embedding, attention and greedy-update kernels read it, so the same
graph serves every step.  `finished` is polled every POLL steps; extra steps past the all-finished point only write
positions that mask_and_clip_seqs discards, so outputs are identical to the reference loop.
"""
from __future__ import annotations

import os

import torch

from . import ops
from .layers import WEIGHTS, Packed, compute_dtype, lens_from_mask, w_

PAGE = 64
POLL = 32


def eligible(decoder, dt) -> bool:
    from .models.kv_caching import CachedTransformerDecoder
    blocks = decoder.decoder_blocks
    if not isinstance(blocks, CachedTransformerDecoder):
        return False
    layer = blocks.layers[0]
    cache_dt = blocks.self_attn_caches[0].k_cache.dtype
    return layer.head_dim == 64 and cache_dt == dt and decoder.hidden_dim % 8 == 0


class GreedyDecodeEngine:
    def __init__(self, decoder, batch, max_len, mem_rows, dt, device):
        self.dec, self.B, self.max_len, self.mem_rows, self.dt, self.device = decoder, batch, max_len, mem_rows, dt, device
        blocks = decoder.decoder_blocks
        self.L, self.D, self.H = len(blocks.layers), decoder.hidden_dim, blocks.layers[0].num_heads
        D, B = self.D, batch
        self.pages_per_seq = (max_len + PAGE - 1) // PAGE
        rows = B * self.pages_per_seq * PAGE
        self.pool = [torch.empty((rows, 2 * D), dtype=dt, device=device) for _ in range(self.L)]
        self.page_table = torch.arange(B * self.pages_per_seq, dtype=torch.int32, device=device).view(B, self.pages_per_seq)
        self.cross = [torch.empty((mem_rows, 2 * D), dtype=dt, device=device) for _ in range(self.L)]
        self.seq_start = torch.zeros(B, dtype=torch.int32, device=device)
        self.kv_lens = torch.zeros(B, dtype=torch.int32, device=device)
        self.step = torch.ones(1, dtype=torch.int32, device=device)
        self.seqs = torch.empty((B, max_len), dtype=torch.long, device=device)
        self.logp = torch.empty((B, max_len), dtype=torch.float32, device=device)
        self.finished = torch.empty(B, dtype=torch.uint8, device=device)
        self.max_mem_len = 0
        self.ws_self = ops.decode_workspace(B, self.H, max_len, device)
        self.ws_cross = None
        # small-M projections: split-K GEMM + fused bias / residual / LayerNorm reduce (ACAI_DECODE_SPLITK=0 switches it off)
        self.use_splitk = os.environ.get("ACAI_DECODE_SPLITK", "1") != "0" and 4 < batch <= 256
        # fp32 residual stream under bf16 autocast (the reference's semantics); in fp32 mode there is only one stream anyway
        self.stream_dt = torch.float32 if os.environ.get("ACAI_DECODE_STREAM", "fp32") != "bf16" else dt
        self.ws_splitk = None
        if self.use_splitk:
            mlp = blocks.layers[0].linear1.out_features
            self.ws_splitk = max((ops.splitk_workspace(batch, D, k, device) for k in (D, mlp)), key=lambda t: t.numel())
        self.graph = None
        self.launches_per_step = 0
        self.logits = None                  # [B, V] logits of the last step (static graph memory once captured)
        self._params_seen = None            # (version, data_ptr) of every decoder parameter + the weight-cache generation
        self._operands_seen = None          # device addresses the captured graph reads its weights from
        # sampling mode (GRPO rollouts): None = greedy; else (top_k, temperature).  The seed is device state so that the
        # captured graph draws a new stream per rollout; `uniforms` (fp32 [B, max_len]) replaces the hash in parity tests.
        self.sampling = None
        self.seed = torch.zeros(1, dtype=torch.int64, device=device)
        self.uniforms = None

    # ---- one decode step (what the graph captures) ------------------------------------------------------------
    def _proj(self, x, weight, bias, residual=None, norm=None, act=ops.ACT_NONE):
        """norm(act(x @ weight.T + bias) + residual): at decode batch sizes (<= 256 rows) one split-K GEMM + one reduce kernel
        with the LayerNorm inside (acai_linear_splitk); otherwise acai_linear_fwd + acai_layernorm_fwd.  With a norm the
        result is the pair (stream row in `stream_dt`, GEMM operand row in `dt`) - one tensor twice when the dtypes agree."""
        dt, sdt = self.dt, self.stream_dt
        two = norm is not None and sdt != dt
        odt = sdt if norm is not None else dt
        if self.use_splitk and dt == torch.bfloat16:
            ln = None if norm is None else (norm.weight, norm.bias, norm.eps)
            y = ops.linear_splitk(x, weight, bias, act=act, residual=residual, ln=ln, out_dtype=odt, workspace=self.ws_splitk,
                                  out2_dtype=dt if two else None)
            if y is not None:
                return y if (two or norm is None) else (y, y)
        y = ops.linear(x, weight, bias, act=act, residual=residual, out_dtype=odt)
        if norm is None:
            return y
        if two:
            return norm(y, out_dtype=sdt, out2_dtype=dt)
        y = norm(y, out_dtype=dt)
        return y, y

    def _step(self):
        dec, dt, D, H = self.dec, self.dt, self.D, self.H
        # x: the residual stream (stream_dt), xb: the same rows as the GEMM operand (dt)
        if self.stream_dt != dt:
            x, xb = ops.decode_embed(self.seqs, self.step, dec.vocab_embedding.weight, dec.pos_embedding, self.stream_dt,
                                     out2_dtype=dt)
        else:
            x = xb = ops.decode_embed(self.seqs, self.step, dec.vocab_embedding.weight, dec.pos_embedding, dt)
        for i, layer in enumerate(dec.decoder_blocks.layers):
            sa, ca = layer.self_attn, layer.multihead_attn
            qkv = ops.linear(xb, w_(sa.in_proj_weight, dt), sa.in_proj_bias, out_dtype=dt)
            a = ops.decode_attn_paged(qkv[:, :D], qkv[:, D:], self.pool[i], H, 64, self.max_len, self.ws_self, dt,
                                      page_table=self.page_table, step=self.step, len_offset=-1)
            x, xb = self._proj(a, w_(sa.out_proj.weight, dt), sa.out_proj.bias, residual=x, norm=layer.norm1)
            qc = self._proj(xb, w_(ca.in_proj_weight, dt, rows=(0, D)), ca.in_proj_bias[:D])
            c = ops.decode_attn_paged(qc, None, self.cross[i], H, 64, self.max_mem_len, self.ws_cross, dt,
                                      seq_start=self.seq_start, kv_lens=self.kv_lens)
            x, xb = self._proj(c, w_(ca.out_proj.weight, dt), ca.out_proj.bias, residual=x, norm=layer.norm2)
            h = ops.linear(xb, w_(layer.linear1.weight, dt), layer.linear1.bias, act=ops.ACT_GELU, out_dtype=dt)
            x, xb = self._proj(h, w_(layer.linear2.weight, dt), layer.linear2.bias, residual=x, norm=layer.norm3)
        if dec.decoder_blocks.norm is not None:
            xb = dec.decoder_blocks.norm(x, out_dtype=dt)
        logits = ops.linear(xb, w_(dec.unembed.weight, dt), dec.unembed.bias, out_dtype=dt)
        self.logits = logits
        if self.sampling is None:
            ops.greedy_update(logits, self.step, self.seqs, self.logp, self.finished, dec.eos_idx)
        else:
            ops.topk_sample_update(logits, self.step, self.seqs, self.logp, self.finished, dec.eos_idx, self.sampling[0],
                                   self.sampling[1], seed=self.seed, uniforms=self.uniforms)
        ops.advance_step(self.step)

    def _reset(self):
        self.seqs.fill_(self.dec.pad_idx)
        self.seqs[:, 0] = self.dec.bos_idx
        self.logp.zero_()
        self.finished.zero_()
        self.step.fill_(1)

    # ---- prefill: cross K/V of the encoder memory straight into the memory cache ----------------------------------
    def prefill(self, mem_packed, mp: Packed):
        dt, D = self.dt, self.D
        if mem_packed.dtype != dt:
            mem_packed = mem_packed.to(dt)
        for i, layer in enumerate(self.dec.decoder_blocks.layers):
            ca = layer.multihead_attn
            ops.linear(mem_packed, w_(ca.in_proj_weight, dt, rows=(D, 3 * D)), ca.in_proj_bias[D:],
                       out=self.cross[i][: mp.total])
        self.seq_start.copy_(mp.cu[:-1], non_blocking=True)
        self.kv_lens.copy_(mp.cu[1:] - mp.cu[:-1], non_blocking=True)
        if mp.max_len > self.max_mem_len or self.ws_cross is None:     # grid of the cross pass grows: re-capture
            self.max_mem_len = max(mp.max_len, self.max_mem_len)
            self.ws_cross = ops.decode_workspace(self.B, self.H, self.max_mem_len, self.device)
            self.graph = None

    def set_sampling(self, sampling, seed=0, uniforms=None):
        """sampling = None (greedy) or (top_k, temperature).  Changing the mode or the uniforms buffer re-captures the
        graph; changing only the seed does not."""
        if sampling is not None:
            sampling = (int(sampling[0]), float(sampling[1]))
        if uniforms is not None:
            if self.uniforms is None or self.uniforms.shape != (self.B, self.max_len):
                self.uniforms = torch.empty((self.B, self.max_len), dtype=torch.float32, device=self.device)
                self.graph = None
            self.uniforms[:, : uniforms.shape[1]].copy_(uniforms)
        elif self.uniforms is not None:
            self.uniforms, self.graph = None, None
        if sampling != self.sampling:
            self.sampling, self.graph = sampling, None
        self.seed.fill_(int(seed) & 0x7FFFFFFFFFFFFFFF)

    # ---- the captured graph holds raw device addresses of the weights: notice when they go stale ---------------------
    def _param_state(self):
        return (WEIGHTS.generation,) + tuple((p._version, p.data_ptr()) for p in self.dec.parameters())

    def _operand_addresses(self):
        """Every weight address `_step` hands to a kernel (the bf16 operand copies `w_` returns, refreshed here if their
        parameter changed, and the fp32 parameters the kernels read directly: biases, LayerNorm rows, embedding tables)."""
        dec, dt, D = self.dec, self.dt, self.D
        ptrs = [p.data_ptr() for p in dec.parameters()]
        for layer in dec.decoder_blocks.layers:
            sa, ca = layer.self_attn, layer.multihead_attn
            for t in (w_(sa.in_proj_weight, dt), w_(sa.out_proj.weight, dt), w_(ca.in_proj_weight, dt, rows=(0, D)),
                      w_(ca.out_proj.weight, dt), w_(layer.linear1.weight, dt), w_(layer.linear2.weight, dt)):
                ptrs.append(t.data_ptr())
        ptrs.append(w_(dec.unembed.weight, dt).data_ptr())
        return tuple(ptrs)

    def _check_weights(self):
        """Drop the graph when a weight it reads has moved (`load_state_dict`, an optimiser step through torch, `.to()`):
        the bf16 operand copy of an updated parameter is a NEW tensor and the old one is freed, so replaying would read
        stale or released memory.  Copies that are refreshed in place (train.FlatTrainer's shadows) keep their address and
        the graph stays valid."""
        state = self._param_state()
        if state == self._params_seen:
            return
        self._params_seen = state
        if self.graph is not None and self._operand_addresses() != self._operands_seen:
            self.graph = None

    def _ensure_graph(self):
        self._check_weights()
        if self.graph is None:
            self._reset()
            n0 = ops.launch_count()
            self._step()                      # eager warm-up: bf16 weight copies, tensor maps, function attributes
            self.launches_per_step = ops.launch_count() - n0
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            self._reset()
            with torch.cuda.graph(g):
                self._step()
            self.graph = g
            self._operands_seen = self._operand_addresses()
            self._params_seen = self._param_state()

    def generate(self, stop_early=True):
        """Run steps t = 1 .. max_len-1 (or until every row has emitted <eos>); returns raw (seqs, log_probs)."""
        self._ensure_graph()
        self._reset()
        for t in range(1, self.max_len):
            self.graph.replay()
            if stop_early and t % POLL == 0 and bool(self.finished.all()):
                break
        return self.seqs, self.logp

    def forced_logits(self, tokens):
        """The captured step driven by GIVEN tokens (B, T): returns (B, T-1, V) logits, entry [:, t-1] being what the greedy
        loop sees at step t after the prefix tokens[:, :t] (input token tokens[:, t-1], position row t, models.py:575-577).
        Same graph, same kernels as `generate`; only the token written back after each step is the caller's.  Parity
        tests compare these step by step with the oracle / the reference's recorded logits."""
        self._ensure_graph()
        self._reset()
        tokens = tokens.to(self.device, torch.long)
        T = min(tokens.shape[1], self.max_len)
        out = []
        for t in range(1, T):
            self.seqs[:, t - 1] = tokens[:, t - 1]
            self.graph.replay()
            out.append(self.logits.float().clone())
        return torch.stack(out, dim=1)


def get_engine(vitomr, decoder, batch, max_len, mem_rows, dt, device):
    cache = vitomr.__dict__.setdefault("_acai_engines", {})
    key = (id(decoder), batch, max_len, dt, str(device))
    eng = cache.get(key)
    if eng is None or eng.mem_rows < mem_rows:
        cache.clear()                          # one engine at a time: the caches are large
        eng = GreedyDecodeEngine(decoder, batch, max_len, max(mem_rows, 1), dt, device)
        cache[key] = eng
    return eng


def greedy_generate(vitomr, img_latent, latent_attention_mask, max_len, stop_early=True, sampling=None, seed=0, uniforms=None):
    """Fast path of ViTOMR.cached_greedy_generate (and, with `sampling` = (top_k, temperature), of
    GRPOViTOMR.cached_forward_rollout_policy); returns None when the decoder does not qualify."""
    dec = vitomr.decoder
    dt = compute_dtype()
    if not img_latent.is_cuda or not eligible(dec, dt) or max_len > dec.max_lmx_seq_len or max_len < 2:
        return None
    B, S = img_latent.shape[0], img_latent.shape[1]
    if B > dec.decoder_blocks.self_attn_caches[0].max_batch_size:
        raise ValueError(f"The current cache has been setup with a max batch size of "
                         f"{dec.decoder_blocks.self_attn_caches[0].max_batch_size}, but found batch size {B}!")
    mem, mp = dec._memory_packed(img_latent, latent_attention_mask, dt)
    eng = get_engine(vitomr, dec, B, max_len, mp.total, dt, img_latent.device)
    eng.prefill(mem, mp)
    eng.set_sampling(sampling, seed, uniforms)
    seqs, logp = eng.generate(stop_early=stop_early)
    return seqs.clone(), logp.clone()
