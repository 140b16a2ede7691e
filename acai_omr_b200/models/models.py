"""Drop-in for `acai_omr/models/models.py` (jsnchon/acai-omr @ 5819ae3) on sm_100a kernels.

Same class names, constructor signatures, forward-path methods, attributes and `state_dict` keys
as the reference; reference line numbers are cited per method.  Internally ragged
batches are never padded: images become packed token rows + `cu_seqlens` straight out of the
patch-embedding kernel, every layer runs on packed rows, and padding only reappears at the API
boundary (`to_padded_tensor` semantics) because callers expect `(B, L_max, E)` + a mask.
"""
from __future__ import annotations

import os
import re

import torch
from torch import nn

from .. import functional as F
from .. import ops
from ..config import LMX_BOS_TOKEN, LMX_EOS_TOKEN, LMX_PAD_TOKEN
from ..layers import (LayerNorm, Linear, Packed, TransformerDecoder, TransformerDecoderLayer, TransformerEncoder,
                      TransformerEncoderLayer, attach_packed, attached_packed, compute_dtype, gelu_mlp_head, lens_from_mask,
                      stream_dtype, to_stream, w_)
from .kv_caching import CachedTransformerDecoder, CachedTransformerDecoderLayer

NUM_CHANNELS = 1  # models.py:11


def _encoder_stack(hidden_dim, num_heads, mlp_dim, dropout, num_layers, final_norm):
    return TransformerEncoder(
        lambda: TransformerEncoderLayer(d_model=hidden_dim, nhead=num_heads, dim_feedforward=mlp_dim, dropout=dropout,
                                        activation="gelu", batch_first=True),
        num_layers=num_layers,
        norm=LayerNorm(hidden_dim, eps=1e-6) if final_norm else None)


class Encoder(nn.Module):
    """models.py:14-96."""
    _interpolate_pe = False

    def __init__(self, patch_size, pe_max_height, pe_max_width, num_layers=12, hidden_dim=768, num_heads=12,
                 mlp_dim=3072, transformer_dropout=0.0):
        super().__init__()
        self.patch_size = patch_size
        self.pe_max_height = pe_max_height
        self.pe_max_width = pe_max_width
        self.hidden_dim = hidden_dim
        self.pos_embedding = nn.Parameter(torch.zeros(self.pe_max_height, self.pe_max_width, self.hidden_dim))
        nn.init.trunc_normal_(self.pos_embedding, std=0.1)
        self.projection = Linear(NUM_CHANNELS * self.patch_size ** 2, self.hidden_dim)
        self.encoder_blocks = _encoder_stack(hidden_dim, num_heads, mlp_dim, transformer_dropout, num_layers, True)

    # ---- packed internals -----------------------------------------------------------------
    def _image_list(self, x):
        imgs = []
        for t in x:  # a bare (1,H,W) tensor iterates as one (H,W) image (vitomr_inference.py:81)
            if t.dim() == 2:
                t = t.unsqueeze(0)
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.float().contiguous()
            imgs.append(t)
        return imgs

    def _grid_check(self, h_p, w_p):
        if not self._interpolate_pe and (h_p > self.pe_max_height or w_p > self.pe_max_width):
            raise ValueError(f"{h_p} x {w_p} image is too large for max positional embedding grid of shape "
                             f"{self.pe_max_height} x {self.pe_max_width}")

    def embed_packed(self, x, out_dtype=None, keep=None):
        """K1: ragged images -> (packed rows [sum L, E], Packed).  `keep`: optional (lens, patch_index) for MAE."""
        imgs = self._image_list(x)
        P = self.patch_size
        lens, dims, oversize = [], [], False
        for t in imgs:
            h_p, w_p = t.shape[-2] // P, t.shape[-1] // P
            self._grid_check(h_p, w_p)
            oversize |= h_p > self.pe_max_height or w_p > self.pe_max_width
            lens.append(h_p * w_p)
            dims.append((h_p, w_p))
        dev = self.pos_embedding.device
        patch_index = None
        if keep is not None:
            lens, patch_index = keep
        pk = Packed(lens, dev)
        geom = F.PatchGeometry(imgs, ops.make_image_descs(imgs, P, dev), pk, P, imgs[0].shape[0], dims, patch_index)
        out = F.patch_embed(geom, self.projection.weight, self.projection.bias, self.pos_embedding,
                            self._interpolate_pe and oversize, out_dtype or compute_dtype())
        return out, pk

    def _stacks(self):
        return [self.encoder_blocks]

    def encode_packed(self, x):
        """Ragged images -> (packed latent rows after all layers + final norm (fp32), Packed)."""
        h, pk = self.embed_packed(x)
        h, hb = to_stream(h)                 # (residual stream rows, GEMM operand rows): one tensor unless bf16 inference
        for s in self._stacks():
            h, hb = s.forward_packed(h, pk, final_dtype=torch.float32 if s.norm is not None else None, xb=hb, return_pair=True)
        return h, pk

    # ---- reference API --------------------------------------------------------------------------
    def batchify(self, x):
        """models.py:36-66: (B, L_m, E) embeddings (+PE) and the (B, L_m) True=padding mask.  Padding rows
        hold projection(0) = the projection bias, as in the reference."""
        h, pk = self.embed_packed(x)
        emb = F.unpack_padded(h, pk.cu, pk.batch, pk.max_len, fill=self.projection.bias)
        return emb, pk.mask(emb.device)

    def create_attention_mask(self, seq_lens, max_len):
        """models.py:70-73."""
        return torch.arange(end=max_len).unsqueeze(0) >= torch.tensor(seq_lens).unsqueeze(1)

    def _pad_fill(self):
        # eval-mode nn.TransformerEncoder returns LayerNorm(0) = norm.bias on padding rows;
        # in train mode the reference leaves arbitrary finite values there; we write zeros.
        last = self._stacks()[-1]
        return last.norm.bias if (last.norm is not None and not self.training) else None

    def forward(self, x):
        """models.py:76-79 -> (x:(B,L_m,E), src_key_padding_mask:(B,L_m) True=padding)."""
        h, pk = self.encode_packed(x)
        out = F.unpack_padded(h, pk.cu, pk.batch, pk.max_len, fill=self._pad_fill())
        mask = pk.mask(out.device)
        attach_packed(out, (h, pk))  # lets ViTOMR feed the decoder without re-packing (dropped if `out` is edited in place)
        return out, mask

    def embed_single_image(self, x):
        """models.py:81-90."""
        h, pk = self.embed_packed([x])
        return h.view(1, pk.total, self.hidden_dim)

    def generate(self, x):
        """models.py:93-96: one (C,H,W) image -> (1, L, E)."""
        h, pk = self.encode_packed([x])
        return h.view(1, pk.total, self.hidden_dim)


class MAEEncoder(Encoder):
    """models.py:100-180: the Encoder with MAE shuffling/masking; only the kept 25 % of the patches are embedded
    (K1 gathers them through a patch index) and encoded."""

    def __init__(self, mask_ratio, patch_size, pe_max_height, pe_max_width, num_layers=12, num_heads=12, hidden_dim=768,
                 mlp_dim=3072):
        super().__init__(patch_size, pe_max_height, pe_max_width, num_layers, hidden_dim, num_heads, mlp_dim,
                         transformer_dropout=0.0)
        self.mask_ratio = mask_ratio

    def mask_indices(self, seq_len, device, noise=None):
        """Index bookkeeping of mask_sequence (models.py:106-125): (ids_keep, ids_restore, seq_mask, len_keep).
        `noise` replaces the per-image torch.rand draw (parity tests inject the reference's draw)."""
        len_keep = int(seq_len * (1 - self.mask_ratio))
        if noise is None:
            noise = torch.rand(seq_len, device=device)
        ids_shuffle = torch.argsort(noise.to(device))
        ids_restore = torch.argsort(ids_shuffle)
        seq_mask = torch.ones(seq_len, device=device, dtype=torch.int)
        seq_mask[:len_keep] = 0
        return ids_shuffle[:len_keep], ids_restore, seq_mask.index_select(0, ids_restore), len_keep

    def mask_sequence(self, t, h_p, w_p, noise=None):
        """models.py:106-125: shuffle + mask one unfolded image t (1, C*P*P, L) -> (t_masked (1, C*P*P, L_keep),
        pos_embed_slice (L_keep, E), unmasked_seq_len, len_keep, seq_mask (L,) int, ids_restore (L,)).  Index bookkeeping
        only (the packed forward path gathers the kept patches inside the patch-embedding kernel instead); `noise`
        replaces the torch.rand draw."""
        unmasked_seq_len = t.shape[-1]
        ids_keep, ids_restore, seq_mask, len_keep = self.mask_indices(unmasked_seq_len, t.device, noise)
        t_masked = t.index_select(dim=-1, index=ids_keep)
        pos_embed_slice = self.pos_embedding[:h_p, :w_p, :].reshape(-1, self.hidden_dim).index_select(dim=0, index=ids_keep.to(
            self.pos_embedding.device))
        return t_masked, pos_embed_slice, unmasked_seq_len, len_keep, seq_mask, ids_restore

    def embed_masked_packed(self, x, noises=None):
        """-> (packed kept-patch embeddings, Packed(kept), info dict for the decoder side)."""
        imgs = self._image_list(x)
        P, dev = self.patch_size, self.pos_embedding.device
        keep, restore, masks, keep_lens, full_lens, dims = [], [], [], [], [], []
        for t in imgs:
            h_p, w_p = t.shape[-2] // P, t.shape[-1] // P
            self._grid_check(h_p, w_p)
            full_lens.append(h_p * w_p); dims.append((h_p, w_p))
            keep_lens.append(int(h_p * w_p * (1 - self.mask_ratio)))
        if noises is None:
            # all images at once: one uniform draw [B, L_max] with the padding pushed behind every real patch, two batched
            # argsorts; the per-image index vectors below are views (the per-image loop cost ~6 tiny kernels per image)
            L_max = max(full_lens)
            lens_t = ops.upload_small(torch.tensor(full_lens, dtype=torch.int64), dev)
            noise = torch.rand((len(imgs), L_max), device=dev)
            noise.masked_fill_(torch.arange(L_max, device=dev)[None, :] >= lens_t[:, None], 2.0)
            ids_shuffle = torch.argsort(noise, dim=1)
            ids_restore = torch.argsort(ids_shuffle, dim=1)
            seq_masks = (ids_restore >= ops.upload_small(torch.tensor(keep_lens, dtype=torch.int64), dev)[:, None]).to(torch.int)
            for i, (L, lk) in enumerate(zip(full_lens, keep_lens)):
                keep.append(ids_shuffle[i, :lk]); restore.append(ids_restore[i, :L]); masks.append(seq_masks[i, :L])
        else:
            for i, L in enumerate(full_lens):
                ids_keep, ids_restore, seq_mask, len_keep = self.mask_indices(L, dev, noises[i])
                assert len_keep == keep_lens[i]
                keep.append(ids_keep); restore.append(ids_restore); masks.append(seq_mask)
        patch_index = torch.cat(keep).to(torch.int32)
        h, pk = self.embed_packed(imgs, keep=(keep_lens, patch_index))
        info = dict(imgs=imgs, kept_seq_lens=keep_lens, unmasked_seq_lens=full_lens, seq_masks=masks, ids_restore=restore,
                    patchified_dims=dims)
        return h, pk, info

    def batchify(self, x, noises=None):
        """models.py:128-173 (padded API; same 8-tuple)."""
        h, pk, info = self.embed_masked_packed(x, noises)
        emb = F.unpack_padded(h, pk.cu, pk.batch, pk.max_len, fill=self.projection.bias)
        full = Packed(info["unmasked_seq_lens"], emb.device)
        return (emb, pk.mask(emb.device), full.mask(emb.device), info["kept_seq_lens"], info["unmasked_seq_lens"],
                torch.nested.as_nested_tensor(info["seq_masks"], layout=torch.jagged),
                torch.nested.as_nested_tensor(info["ids_restore"], layout=torch.jagged), info["patchified_dims"])

    def encode_masked_packed(self, x, noises=None):
        h, pk, info = self.embed_masked_packed(x, noises)
        h = self.encoder_blocks.forward_packed(h, pk, final_dtype=torch.float32)
        return h, pk, info

    def forward(self, x, noises=None):
        """models.py:176-180."""
        h, pk, info = self.encode_masked_packed(x, noises)
        out = F.unpack_padded(h, pk.cu, pk.batch, pk.max_len, fill=self._pad_fill())
        attach_packed(out, (h, pk, info))
        full = Packed(info["unmasked_seq_lens"], out.device)
        return (out, full.mask(out.device), info["kept_seq_lens"], info["unmasked_seq_lens"],
                torch.nested.as_nested_tensor(info["seq_masks"], layout=torch.jagged),
                torch.nested.as_nested_tensor(info["ids_restore"], layout=torch.jagged), info["patchified_dims"])


class MAEDecoder(nn.Module):
    """models.py:182-195."""

    def __init__(self, num_layers=8, hidden_dim=512, num_heads=16, mlp_dim=3072, transformer_dropout=0.0):
        super().__init__()
        self.decoder_blocks = _encoder_stack(hidden_dim, num_heads, mlp_dim, transformer_dropout, num_layers, True)

    def forward(self, x, attention_mask):
        return self.decoder_blocks(x, src_key_padding_mask=attention_mask)


class MAE(nn.Module):
    """models.py:197-269."""

    def __init__(self, mask_ratio, patch_size, pe_max_height, pe_max_width, encoder_hidden_dim=768, decoder_hidden_dim=512,
                 encoder_kwargs={}, decoder_kwargs={}):
        super().__init__()
        self.patch_size = patch_size
        self.encoder = MAEEncoder(mask_ratio, self.patch_size, pe_max_height, pe_max_width, hidden_dim=encoder_hidden_dim,
                                  **encoder_kwargs)
        self.decoder_hidden_dim = decoder_hidden_dim
        self.decoder = MAEDecoder(hidden_dim=self.decoder_hidden_dim, **decoder_kwargs)
        self.decoder_embed = Linear(encoder_hidden_dim, self.decoder_hidden_dim)
        self.decoder_unembed = Linear(self.decoder_hidden_dim, NUM_CHANNELS * patch_size ** 2)
        self.mask_token = nn.Parameter(torch.zeros(1, 1, self.decoder_hidden_dim))
        self.decoder_pos_embedding = nn.Parameter(torch.zeros(pe_max_height, pe_max_width, self.decoder_hidden_dim))
        nn.init.trunc_normal_(self.mask_token, std=0.1)
        nn.init.trunc_normal_(self.decoder_pos_embedding, std=0.1)

    def _restore_packed(self, latent_packed, pk_keep, info, out_dtype):
        dev = latent_packed.device
        full = Packed(info["unmasked_seq_lens"], dev)
        ids = torch.cat([r.to(dev) for r in info["ids_restore"]]).to(torch.int64).contiguous()
        wps = ops.upload_small(torch.tensor([w for _, w in info["patchified_dims"]], dtype=torch.int32), dev)
        rows = F.mae_restore(latent_packed, pk_keep.cu, full.cu, ids, wps, full.total, self.mask_token,
                             self.decoder_pos_embedding, out_dtype)
        return rows, full

    def prepare_for_decoder(self, latent, kept_seq_lens, unmasked_seq_lens, batch_ids_restore, patchified_dims):
        """models.py:219-241 (padded API): latent (N, L_keep_max, D) -> (N, L_max, D) with mask tokens, unshuffled, + PE."""
        dev = latent.device
        pk = Packed(kept_seq_lens, dev)
        restore = batch_ids_restore.unbind() if hasattr(batch_ids_restore, "unbind") else list(batch_ids_restore)
        info = dict(unmasked_seq_lens=unmasked_seq_lens, ids_restore=list(restore), patchified_dims=patchified_dims)
        packed = F.pack_padded(latent, pk.cu, pk.total)
        rows, full = self._restore_packed(packed, pk, info, latent.dtype)
        return F.unpack_padded(rows, full.cu, full.batch, full.max_len)

    def forward(self, batch, noises=None):
        """models.py:249-269 -> (pred (N, L_m, C*P*P), loss_mask (N, L_m) bool, target (N, L_m, C*P*P))."""
        x, y = zip(*batch)
        dt = compute_dtype()
        h, pk, info = self.encoder.encode_masked_packed(list(x), noises)
        lat = self.decoder_embed(h, out_dtype=dt)
        rows, full = self._restore_packed(lat, pk, info, dt)
        hid = self.decoder.decoder_blocks.forward_packed(rows, full, final_dtype=torch.float32)
        pred = self.decoder_unembed(hid, out_dtype=dt)
        dev = pred.device
        P = self.patch_size
        tgts = self.encoder._image_list(list(y))
        tgt_rows = ops.patchify(tgts, ops.make_image_descs(tgts, P, dev), full.cu, full.total, P, tgts[0].shape[0])
        pred_p = F.unpack_padded(pred, full.cu, full.batch, full.max_len)
        target_p = ops.unpack_padded(tgt_rows, full.cu, full.batch, full.max_len)
        loss_mask = torch.zeros((full.batch, full.max_len), dtype=torch.bool, device=dev)
        for i, m in enumerate(info["seq_masks"]):
            loss_mask[i, : m.shape[0]] = m.bool()
        return pred_p, loss_mask, target_p


class MAELoss(nn.Module):
    """models.py:271-288."""

    def forward(self, pred, loss_mask, target):
        return F.mae_loss(pred, loss_mask, target)[0]


class OMREncoder(Encoder):
    """models.py:290-332: as Encoder, but oversize images get a bilinearly resized PE grid."""
    _interpolate_pe = True

    def interpolate_pe(self, h_p, w_p):
        """models.py:291-302.  Evaluated by the patch-embedding kernel's 4-tap sampler on a zero image."""
        P = self.patch_size
        zero = torch.zeros(NUM_CHANNELS, h_p * P, w_p * P, device=self.pos_embedding.device)
        pk = Packed([h_p * w_p], zero.device)
        descs = ops.make_image_descs([zero], P, zero.device)
        pe = ops.patch_embed([zero], descs, pk.cu, pk.total, P, NUM_CHANNELS, self.projection.weight, None,
                             self.pos_embedding, True, torch.float32)
        return pe.view(h_p, w_p, self.hidden_dim)


class FineTuneOMREncoder(OMREncoder):
    """models.py:334-376."""

    def __init__(self, patch_size, pe_max_height, pe_max_width, fine_tune_depth, num_layers=12, hidden_dim=768,
                 num_heads=12, mlp_dim=3072, transformer_dropout=0.05):
        super().__init__(patch_size, pe_max_height, pe_max_width, num_layers, hidden_dim, num_heads, mlp_dim)
        assert fine_tune_depth > 0, "If using FineTuneOMREncoder, fine-tune depth should be at least 1"
        del self.encoder_blocks
        self.fine_tune_depth = fine_tune_depth
        self.num_layers = num_layers
        self.num_frozen_layers = self.num_layers - self.fine_tune_depth
        self.superclass_kwargs = {"num_heads": num_heads, "mlp_dim": mlp_dim, "transformer_dropout": transformer_dropout}
        if self.num_frozen_layers == 0:
            self.frozen_blocks = None
        else:
            self.frozen_blocks = _encoder_stack(hidden_dim, num_heads, mlp_dim, 0.0, self.num_frozen_layers, False)
        self.fine_tune_blocks = _encoder_stack(hidden_dim, num_heads, mlp_dim, transformer_dropout,
                                               self.fine_tune_depth, True)

    def _stacks(self):
        return ([self.frozen_blocks] if self.frozen_blocks is not None else []) + [self.fine_tune_blocks]


class TransitionHead(nn.Sequential):
    """nn.Sequential(Linear, GELU, Dropout, Linear) with the reference's child indices ("0", "3";
    models.py:655-660); forward fuses the GELU into the first projection's epilogue."""

    def forward(self, x):
        p = self[2].p if self.training else 0.0
        packed = attached_packed(x)
        if packed is None:
            return gelu_mlp_head(x, self[0], self[3], dropout_p=p)
        # came straight from the encoder: run on the packed rows of valid patches only.  The reference applies the head to
        # the padding rows too; in eval mode those all hold the same vector (the encoder's final-norm bias), so the head
        # maps them to ONE row, computed once here and used as the fill.  (Train mode: the reference's padding rows are
        # unspecified finite values and nothing downstream reads them; they are zero here.)
        h, pk = packed
        hp = gelu_mlp_head(h, self[0], self[3], dropout_p=p)
        fill = None
        if not self.training:
            short = next((b for b, n in enumerate(pk.lens) if n < x.shape[1]), None)
            if short is not None:
                fill = gelu_mlp_head(x[short, pk.lens[short]].reshape(1, -1), self[0], self[3]).reshape(-1).float()
        out = F.unpack_padded(hp, pk.cu, pk.batch, x.shape[1], fill=fill)
        attach_packed(out, (hp, pk))
        return out


class OMRDecoder(nn.Module):
    """models.py:378-528."""

    def __init__(self, max_lmx_seq_len, lmx_vocab_path, num_layers=10, hidden_dim=1024, num_heads=16, mlp_dim=4096,
                 transformer_dropout=0.1, use_caching=False, max_batch_size=None, cache_dtype=None):
        super().__init__()
        self.max_lmx_seq_len = max_lmx_seq_len
        self.lmx_vocab_path = lmx_vocab_path
        self.num_layers = num_layers
        self.hidden_dim = hidden_dim
        self.num_heads = num_heads
        self.head_dim = hidden_dim / num_heads  # a float in the reference too (models.py:388)
        self.mlp_dim = mlp_dim
        self.transformer_dropout = transformer_dropout
        with open(lmx_vocab_path, "r") as f:
            tokens = [line.strip() for line in f if line.strip()]
        self.tokens_to_idxs = {token: i for i, token in enumerate(tokens)}
        self.idxs_to_tokens = {i: token for i, token in enumerate(tokens)}
        self.pad_idx = self.tokens_to_idxs[LMX_PAD_TOKEN]
        self.bos_idx = self.tokens_to_idxs[LMX_BOS_TOKEN]
        self.eos_idx = self.tokens_to_idxs[LMX_EOS_TOKEN]
        self.vocab_size = len(tokens)
        self.vocab_embedding = nn.Embedding(self.vocab_size, self.hidden_dim, padding_idx=self.pad_idx)
        self.pos_embedding = nn.Parameter(torch.zeros(self.max_lmx_seq_len, self.hidden_dim))
        nn.init.trunc_normal_(self.pos_embedding, std=0.1)
        layer_kwargs = dict(d_model=self.hidden_dim, nhead=num_heads, dim_feedforward=mlp_dim,
                            dropout=transformer_dropout, activation="gelu", batch_first=True)
        if use_caching:
            self.decoder_blocks = CachedTransformerDecoder(
                decoder_layer=CachedTransformerDecoderLayer(**layer_kwargs), num_layers=num_layers,
                max_batch_size=max_batch_size, max_decoder_seq_len=max_lmx_seq_len, cache_dtype=cache_dtype,
                norm=LayerNorm(self.hidden_dim, eps=1e-6))
        else:
            self.decoder_blocks = TransformerDecoder(lambda: TransformerDecoderLayer(**layer_kwargs),
                                                     num_layers=num_layers, norm=LayerNorm(self.hidden_dim, eps=1e-6))
        self.unembed = Linear(self.hidden_dim, self.vocab_size)

    def to_cached_version(self, max_batch_size, cache_dtype):
        """models.py:431-443."""
        return OMRDecoder(self.max_lmx_seq_len, self.lmx_vocab_path, self.num_layers, self.hidden_dim, self.num_heads,
                          self.mlp_dim, self.transformer_dropout, use_caching=True, max_batch_size=max_batch_size,
                          cache_dtype=cache_dtype)

    # ---- teacher-forced / uncached -----------------------------------------------------------
    def _memory_packed(self, img_latent, latent_attention_mask, dt):
        packed = attached_packed(img_latent)
        B, S = img_latent.shape[0], img_latent.shape[1]
        lens = lens_from_mask(latent_attention_mask, B, S)
        if packed is not None and packed[1].lens == lens:
            mem, mp = packed
            return (mem if mem.dtype == dt else mem.to(dt)), mp
        mp = Packed(lens, img_latent.device)
        return F.pack_padded(img_latent, mp.cu, mp.total, out_dtype=dt), mp

    def _decode_packed(self, input_seqs, img_latent, tgt_lens, latent_attention_mask, token_idxs_input, checkpoint_grads=False):
        B, T = input_seqs.shape[0], input_seqs.shape[1]
        dt = compute_dtype()
        dev = self.pos_embedding.device
        tp = Packed(tgt_lens, dev)
        stream = None
        if token_idxs_input:
            sdt = stream_dtype(dt)           # fp32 embedding + position sum when the residual stream is fp32
            x = F.embed_tokens(input_seqs, self.vocab_embedding.weight, self.pos_embedding, sdt,
                               padding_idx=self.vocab_embedding.padding_idx)  # (B,T,D)
            x = F.pack_padded(x, tp.cu, tp.total)
            if sdt != dt:
                stream, x = x, x.to(dt)
        else:   # embedding-space inputs (scheduled sampling, models.py:463-470): add the position table, then pack
            x = input_seqs if input_seqs.dtype == dt else input_seqs.to(dt)
            x = F.mix_rows(x, pos=self.pos_embedding, period=T)
            x = F.pack_padded(x, tp.cu, tp.total)
        mem, mp = self._memory_packed(img_latent, latent_attention_mask, dt)
        h = self.decoder_blocks.forward_packed(x, tp, mem, mp, causal=True, stream=stream, checkpoint=checkpoint_grads)
        logits = F.linear(h, self.unembed.weight, self.unembed.bias, out_dtype=dt)
        return F.unpack_padded(logits, tp.cu, B, T)

    def forward(self, input_seqs, img_latent, lmx_attention_mask, latent_attention_mask, token_idxs_input=True,
                checkpoint_grads=False):
        """models.py:445-483 -> (B, T, vocab).  Rows of padded target positions come back as zeros (the
        reference leaves unspecified finite values there; the loss ignores them)."""
        T = input_seqs.shape[1]
        if T > self.max_lmx_seq_len:
            raise ValueError(f"{T} long lmx sequence length is too long for max sequence length of {self.max_lmx_seq_len}")
        lens = lens_from_mask(lmx_attention_mask, input_seqs.shape[0], T)
        return self._decode_packed(input_seqs, img_latent, lens, latent_attention_mask, token_idxs_input,
                                   checkpoint_grads=checkpoint_grads)

    def generate(self, input_seqs, img_latent, latent_attention_mask=None):
        """models.py:487-500."""
        seq_len = input_seqs.shape[1]
        if seq_len > self.max_lmx_seq_len:
            raise ValueError(f"{seq_len} long lmx sequence length is too long for max sequence length of {self.max_lmx_seq_len}")
        return self._decode_packed(input_seqs, img_latent, [seq_len] * input_seqs.shape[0], latent_attention_mask, True)

    # ---- KV-cached ------------------------------------------------------------------------------
    def prepare_caches(self, encoder_memory):
        """models.py:503-507."""
        if not isinstance(self.decoder_blocks, CachedTransformerDecoder):
            raise RuntimeError("Trying to use cached inference pathway with an uncached TransformerDecoder instance")
        self.decoder_blocks.prepare_caches(encoder_memory)

    def cached_generate(self, token_t, time_step, latent_attention_mask=None):
        """models.py:518-528: token_t (B,1) -> next-token logits (B,1,V)."""
        if time_step >= self.max_lmx_seq_len:
            raise RuntimeError(f"{time_step + 1} decoding steps is too long for max sequence length of {self.max_lmx_seq_len}")
        dt = compute_dtype()
        emb = ops.embed_tokens(token_t, self.vocab_embedding.weight, self.pos_embedding, stream_dtype(dt), pos_offset=time_step)
        h = self.decoder_blocks.cached_generate(emb, memory_key_padding_mask=latent_attention_mask)
        return ops.linear(h, w_(self.unembed.weight, dt), self.unembed.bias, out_dtype=dt)


def batchify_and_split_lmx_seqs(lmx_seqs, pad_idx, device):
    """models.py:531-540 (host-side index bookkeeping; no arithmetic)."""
    lens = [int(s.shape[0]) for s in lmx_seqs]
    T = max(lens)
    src_dev = lmx_seqs[0].device
    padded = torch.full((len(lmx_seqs), T), pad_idx, dtype=torch.long, device=src_dev)
    if src_dev.type == "cuda" and len(lmx_seqs) > 1:
        # one concatenation + one scatter instead of a copy kernel per sequence; the (row, column) tables are built on
        # the host from the lengths and uploaded asynchronously
        lens_t = torch.tensor(lens)
        rows = torch.repeat_interleave(torch.arange(len(lens)), lens_t)
        cols = torch.arange(int(lens_t.sum())) - torch.repeat_interleave(torch.cumsum(lens_t, 0) - lens_t, lens_t)
        padded[ops.upload_small(rows, src_dev), ops.upload_small(cols, src_dev)] = torch.cat([s.to(torch.long) for s in lmx_seqs])
    else:
        for i, s in enumerate(lmx_seqs):
            padded[i, : s.shape[0]] = s
    input_seqs, target_seqs = padded[:, :-1], padded[:, 1:]
    lmx_attention_mask = (input_seqs == pad_idx).to(device)
    return input_seqs, target_seqs, lmx_attention_mask


class ViTOMR(nn.Module):
    """models.py:542-647."""

    def __init__(self, encoder, transition_head, decoder):
        super().__init__()
        self.encoder = encoder
        self.transition_head = transition_head
        self.decoder = decoder

    def create_inference_mask(self, seqs):
        """models.py:550-559 (integer bookkeeping on the finished sequences)."""
        eos_mask = seqs == self.decoder.eos_idx
        seen_eos_mask = eos_mask.int().cumsum(dim=-1)
        first_eos_mask = eos_mask & (seen_eos_mask == 1)
        return (seen_eos_mask == 0) | first_eos_mask

    def cached_set_up_inference(self, img_latent, max_len):
        """models.py:562-573; `finished` lives on the device here (no per-step host sync)."""
        device = img_latent.device
        self.decoder.prepare_caches(img_latent)
        num_seqs = img_latent.shape[0]
        seqs = torch.full([num_seqs, max_len], fill_value=self.decoder.pad_idx, dtype=torch.long, device=device)
        seqs[:, 0] = self.decoder.bos_idx
        seq_log_probs = torch.zeros_like(seqs, dtype=torch.float)
        finished_seqs = torch.zeros([num_seqs], dtype=torch.bool, device=device)
        return seqs, seq_log_probs, finished_seqs

    def cached_get_next_token(self, seqs, t, latent_attention_mask):
        """models.py:575-583, including `time_step = t` for the token at index t-1."""
        logits = self.decoder.cached_generate(seqs[:, t - 1].unsqueeze(1), t, latent_attention_mask).squeeze(1)
        return ops.argmax_logprob(logits)

    def mask_and_clip_seqs(self, seqs, seq_log_probs):
        """models.py:585-596."""
        seq_mask = self.create_inference_mask(seqs)
        seqs = seqs.masked_fill(~seq_mask, self.decoder.pad_idx)
        seq_log_probs = seq_log_probs.masked_fill(~seq_mask, 0.0)
        max_seq_len = int(torch.max(seq_mask.sum(dim=-1)))
        return seqs[:, :max_seq_len], seq_log_probs[:, :max_seq_len], seq_mask[:, :max_seq_len]

    EOS_POLL_INTERVAL = 16

    def cached_greedy_generate(self, img_latent, latent_attention_mask=None, max_len=1536):
        """models.py:600-615.  The all-finished test is polled every EOS_POLL_INTERVAL steps instead of
        every step; the extra steps only write positions after every row's first <eos>, which
        mask_and_clip_seqs discards, so the returned tensors are identical."""
        from ..decode_engine import greedy_generate
        fast = greedy_generate(self, img_latent, latent_attention_mask, max_len)   # CUDA-graphed paged-KV engine
        if fast is not None:
            return self.mask_and_clip_seqs(*fast)
        # generic path (any head_dim / cache dtype): the module-level cached API, one launch per op
        seqs, seq_log_probs, finished = self.cached_set_up_inference(img_latent, max_len)
        for t in range(1, max_len):
            nxt, lp = self.cached_get_next_token(seqs, t, latent_attention_mask)
            seqs[:, t] = nxt
            seq_log_probs[:, t] = lp
            finished |= nxt == self.decoder.eos_idx
            if t % self.EOS_POLL_INTERVAL == 0 and bool(finished.all()):
                break
        return self.mask_and_clip_seqs(seqs, seq_log_probs)

    def streamed_cached_greedy_generate(self, img_latent, latent_attention_mask=None, max_len=1536, flush_interval=25):
        """models.py:625-647 (event dicts use the reference's InferenceEvent string values)."""
        if img_latent.shape[0] != 1:
            raise ValueError("Streamed generation only supports single image batches")
        seqs, seq_log_probs, finished = self.cached_set_up_inference(img_latent, max_len)
        token_buffer = torch.full([seqs.shape[0], flush_interval], fill_value=self.decoder.pad_idx, dtype=torch.int,
                                  device=seqs.device)
        for t in range(1, max_len):
            nxt, lp = self.cached_get_next_token(seqs, t, latent_attention_mask)
            seqs[:, t] = nxt
            seq_log_probs[:, t] = lp
            token_buffer[:, (t - 1) % flush_interval] = nxt
            finished |= nxt == self.decoder.eos_idx
            if bool(finished.all()):
                break
            if t % flush_interval == 0:
                yield {"type": "step", "payload": {"tokens": token_buffer}}
        seqs, seq_log_probs, seq_mask = self.mask_and_clip_seqs(seqs, seq_log_probs)
        yield {"type": "inference_finish", "payload": {"sequence": seqs, "log_probs": seq_log_probs, "mask": seq_mask}}


class TeacherForcedViTOMR(ViTOMR):
    """models.py:649-781."""

    def __init__(self, omr_encoder, pretrained_mae_state_dict, omr_decoder, transition_head_dim=4096,
                 transition_head_dropout=0.05):
        encoder, decoder = omr_encoder, omr_decoder
        transition_head = TransitionHead(
            Linear(encoder.hidden_dim, transition_head_dim), nn.GELU(), nn.Dropout(transition_head_dropout),
            Linear(transition_head_dim, decoder.hidden_dim))
        super().__init__(encoder, transition_head, decoder)
        if pretrained_mae_state_dict:
            encoder.load_state_dict(self.create_omr_encoder_state_dict_from_mae(pretrained_mae_state_dict))
        if isinstance(self.encoder, FineTuneOMREncoder) and self.encoder.frozen_blocks:
            for param in self.encoder.frozen_blocks.parameters():
                param.requires_grad = False
            for param in self.encoder.projection.parameters():
                param.requires_grad = False
            self.encoder.pos_embedding.requires_grad = False
        elif isinstance(self.encoder, OMREncoder) and not isinstance(self.encoder, FineTuneOMREncoder):
            for param in self.encoder.parameters():
                param.requires_grad = False

    def create_omr_encoder_state_dict_from_mae(self, pretrained_mae_state_dict):
        """models.py:679-713: MAE `encoder.encoder_blocks.*` keys -> frozen/fine-tune block keys."""
        sd = {k[len("encoder."):]: v for k, v in pretrained_mae_state_dict.items() if k.startswith("encoder.")}
        if not isinstance(self.encoder, FineTuneOMREncoder):
            return sd
        out = {}
        threshold = self.encoder.num_layers - self.encoder.fine_tune_depth
        layer_num = re.compile(r"(?:\w|\.)+?(\d+)(?:\w|\.)+")
        for key, value in sd.items():
            m = layer_num.match(key)
            if m:
                n = int(m.group(1))
                if n < threshold:
                    new = key.replace("encoder_blocks", "frozen_blocks")
                else:
                    new = key.replace("encoder_blocks", "fine_tune_blocks").replace(f"layers.{n}", f"layers.{n - threshold}")
                out[new] = value
            elif key in ("encoder_blocks.norm.weight", "encoder_blocks.norm.bias"):
                out[key.replace("encoder_blocks", "fine_tune_blocks")] = value
            else:
                out[key] = value
        return out

    def forward(self, x):
        """models.py:722-736: list[(img, lmx_seq)] -> (pred (B,T,V), target_seqs (B,T))."""
        imgs, lmx_seqs = zip(*x)
        img_latent, latent_attention_mask = self.encoder(imgs)
        hook = getattr(self, "_acai_latent_hook", None)       # train.teacher_forced_step: gradient all-reduce overlap
        if hook is not None:
            # the gradient reaches the encoder through the PACKED rows the transition head consumes (the padded tensor is a
            # dead end of the tape when its packed note is used), so that is where the hook belongs
            packed = attached_packed(img_latent)
            hook(packed[0] if packed is not None else img_latent)
        img_latent = self.transition_head(img_latent)
        input_seqs, target_seqs, lmx_attention_mask = batchify_and_split_lmx_seqs(lmx_seqs, self.decoder.pad_idx,
                                                                                  img_latent.device)
        T = input_seqs.shape[1]
        lmx_attention_mask._acai_lens = [min(int(s.shape[0]), T) for s in lmx_seqs]
        pred = self.decoder(input_seqs, img_latent, lmx_attention_mask, latent_attention_mask)
        return pred, target_seqs

    def generate(self, img_latent, seqs):
        """models.py:749-757: next-token log-distribution (num_seqs, vocab) for `seqs` (num_seqs, T) that all continue the
        one image whose latent (1, S, E) is given (beam-search style expansion)."""
        img_latent = img_latent.expand(seqs.shape[0], -1, -1)
        logits = self.decoder.generate(seqs, img_latent)
        return torch.log_softmax(logits[:, -1, :].float(), dim=-1).to(logits.dtype)

    def create_fine_tune_param_groups(self, base_lr, fine_tune_base_lr, fine_tune_decay_factor):
        """models.py:761-781."""
        param_groups = [{"params": self.decoder.parameters(), "lr": base_lr},
                        {"params": self.transition_head.parameters(), "lr": base_lr}]
        layer_lrs = []
        for i, layer in enumerate(reversed(self.encoder.fine_tune_blocks.layers)):
            layer_lr = fine_tune_base_lr * (fine_tune_decay_factor ** i)
            param_groups.append({"params": layer.parameters(), "lr": layer_lr})
            layer_lrs.append(layer_lr)
        min_layer_lr = layer_lrs[-1]
        param_groups.append({"params": self.encoder.fine_tune_blocks.norm.parameters(), "lr": fine_tune_base_lr})
        param_groups.append({"params": (p for p in [self.encoder.pos_embedding]), "lr": min_layer_lr})
        param_groups.append({"params": self.encoder.projection.parameters(), "lr": min_layer_lr})
        return param_groups, layer_lrs


class ScheduledSamplingViTOMR(TeacherForcedViTOMR):
    """models.py:798-838: a second decoder pass on inputs mixed from gold embeddings and the first pass' expected
    embeddings (Gumbel-softmax of its logits, right-shifted)."""

    def sample_and_mix_seqs(self, teacher_forcing_prob, tf_input_seqs, tf_pred_logits, sample_tau, use_hard_sampling, device,
                            sample_rand=None, gumbel_noise=None):
        """models.py:801-817.  `sample_rand` (B,T) replaces the torch.rand draw and `gumbel_noise` (B,T,V) the Gumbel(0,1)
        draw of F.gumbel_softmax (parity tests inject the reference's draws)."""
        B, T = tf_input_seqs.shape
        dec = self.decoder
        sampling_ratio = 1 - teacher_forcing_prob
        if sample_rand is None:
            sample_rand = torch.rand((B, T), device=device)
        sample_mask = sample_rand.to(device) < sampling_ratio
        if gumbel_noise is None:
            gumbel_noise = -torch.empty(tf_pred_logits.shape, dtype=torch.float32, device=device).exponential_().log()
        dt = tf_pred_logits.dtype
        ids = tf_input_seqs.to(device)
        gold = F.embed_rows(ids, dec.vocab_embedding.weight, dt, padding_idx=dec.vocab_embedding.padding_idx)   # (B,T,E)
        distrs = F.gumbel_softmax(tf_pred_logits, gumbel_noise.to(device), sample_tau, use_hard_sampling)        # (B,T,V)
        expected = F.expected_embedding(distrs, dec.vocab_embedding.weight)                                       # (B,T,E)
        # position t takes the expectation predicted at t-1 (right shift); position 0 always keeps the gold <bos> row
        t = torch.arange(T, device=device)
        row = torch.arange(B * T, device=device).view(B, T)
        sel = torch.where(sample_mask & (t > 0)[None, :], row - 1, torch.full_like(row, -1))
        return F.mix_rows(gold, a=expected, sel=sel)

    def forward_train(self, x, teacher_forcing_prob, sample_tau, use_hard_sampling, sample_rand=None, gumbel_noise=None):
        """models.py:819-834 -> (pred (B,T,V), target_seqs (B,T))."""
        imgs, lmx_seqs = zip(*x)
        img_latent, latent_attention_mask = self.encoder(imgs)
        img_latent = self.transition_head(img_latent)
        device = img_latent.device
        tf_input_seqs, target_seqs, lmx_attention_mask = batchify_and_split_lmx_seqs(lmx_seqs, self.decoder.pad_idx, device)
        T = tf_input_seqs.shape[1]
        lmx_attention_mask._acai_lens = [min(int(s.shape[0]), T) for s in lmx_seqs]
        tf_pred_logits = self.decoder(tf_input_seqs, img_latent, lmx_attention_mask, latent_attention_mask)
        mixed = self.sample_and_mix_seqs(teacher_forcing_prob, tf_input_seqs, tf_pred_logits, sample_tau, use_hard_sampling,
                                         device, sample_rand, gumbel_noise)
        pred = self.decoder(mixed, img_latent, lmx_attention_mask, latent_attention_mask, token_idxs_input=False)
        return pred, target_seqs

    def forward_eval(self, x):
        """models.py:836-838."""
        return super().forward(x)


class GRPOViTOMR(ViTOMR):
    """models.py:840-1049: the ViTOMR prepared for GRPO - a full (non-fine-tune) encoder and the transition head frozen
    with their dropout off, weights taken from a TeacherForcedViTOMR state dict, and the rollout policy: top-k filtering,
    temperature softmax and one multinomial draw per decode step.  The cached rollout runs on the CUDA-graphed decode
    engine with the draw done on the device (`acai_topk_sample_update`), so a rollout has no per-step host round trip."""

    def __init__(self, encoder, transition_head, decoder, teacher_forced_state_dict):
        super().__init__(encoder, transition_head, decoder)
        if isinstance(self.encoder, FineTuneOMREncoder):
            teacher_forced_state_dict = self.convert_teacher_forced_state_dict(teacher_forced_state_dict,
                                                                               encoder.num_frozen_layers)
            e = self.encoder
            self.encoder = OMREncoder(e.patch_size, e.pe_max_height, e.pe_max_width, e.num_layers, e.hidden_dim,
                                      **encoder.superclass_kwargs)
        self.load_state_dict(teacher_forced_state_dict)
        self.freeze_component(self.encoder)
        self.freeze_component(self.transition_head)

    def freeze_component(self, component):
        """models.py:855-860: no gradients, no dropout (nn.Dropout children and this package's `dropout_p` attributes)."""
        for param in component.parameters():
            param.requires_grad = False
        for child in component.modules():
            if isinstance(child, nn.Dropout):
                child.p = 0.0
            if hasattr(child, "dropout_p"):
                child.dropout_p = 0.0

    def convert_teacher_forced_state_dict(self, teacher_forced_state_dict, num_frozen_layers):
        """models.py:862-882: `frozen_blocks.layers.N` -> `encoder_blocks.layers.N`, `fine_tune_blocks.layers.N` ->
        `encoder_blocks.layers.(N + num_frozen_layers)`, the stack's final norm just renamed."""
        out = {}
        for name, value in teacher_forced_state_dict.items():
            if "frozen_blocks" in name:
                name = name.replace("frozen_blocks", "encoder_blocks")
            elif "fine_tune_blocks" in name:
                m = re.search(r"fine_tune_blocks\.layers\.(\d+)\.", name)
                name = name.replace("fine_tune_blocks", "encoder_blocks")
                if m is not None:
                    n = int(m.group(1))
                    name = name.replace(f"layers.{n}.", f"layers.{n + num_frozen_layers}.", 1)
            out[name] = value
        return out

    def expand_img_latent_for_rollout(self, img_latent, latent_attention_mask, group_size):
        """models.py:885-894: every latent (and its mask row) repeated `group_size` times, groups contiguous."""
        B, T, E = img_latent.shape
        latent = img_latent.unsqueeze(1).expand(B, group_size, T, E).flatten(0, 1)
        mask = latent_attention_mask.unsqueeze(1).expand(B, group_size, T).flatten(0, 1)
        lens = getattr(latent_attention_mask, "_acai_lens", None)
        if lens is not None:                                   # keeps the no-sync re-packing path of the decoder alive
            mask = mask.contiguous()
            mask._acai_lens = [v for v in lens for _ in range(group_size)]
        return latent, mask

    def _finish_rollouts(self, rollouts, rollout_log_probs):
        """models.py:1031-1049: mask after the first <eos>, clip to the longest rollout."""
        rollout_mask = self.create_inference_mask(rollouts)
        rollouts = rollouts.masked_fill(~rollout_mask, self.decoder.pad_idx)
        max_rollout_len = int(torch.max(rollout_mask.sum(dim=-1)))
        rollouts, rollout_log_probs = rollouts[:, :max_rollout_len], rollout_log_probs[:, :max_rollout_len]
        rollout_mask = rollout_mask[:, :max_rollout_len]
        return rollouts, rollout_log_probs.masked_fill(~rollout_mask, 0.0), rollout_mask

    def cached_forward_rollout_policy(self, img_latent, latent_attention_mask, max_actions=768, top_k=50, temperature=1.2,
                                      seed=None, uniforms=None):
        """models.py:991-1049.  `seed` (int) makes a rollout replayable; `uniforms` (fp32 [R, max_actions], entry [r, t]
        drives the draw of token t) replaces the random stream entirely (parity tests)."""
        from ..decode_engine import greedy_generate
        if seed is None:
            seed = ops.DropoutState.next_seed()
        if uniforms is not None:
            uniforms = uniforms.to(device=img_latent.device, dtype=torch.float32).contiguous()
        fast = greedy_generate(self, img_latent, latent_attention_mask, max_actions, sampling=(top_k, temperature), seed=seed,
                               uniforms=uniforms)
        if fast is not None:
            return self._finish_rollouts(*fast)
        # generic path (any head_dim / cache dtype): module-level cached API, the draw still on the device
        rollouts, log_probs, finished = self.cached_set_up_inference(img_latent, max_actions)
        finished = finished.to(torch.uint8)
        seed_t = torch.full((1,), int(seed) & 0x7FFFFFFFFFFFFFFF, dtype=torch.int64, device=img_latent.device)
        step = torch.zeros(1, dtype=torch.int32, device=img_latent.device)
        for t in range(1, max_actions):
            logits = self.decoder.cached_generate(rollouts[:, t - 1].unsqueeze(1), t, latent_attention_mask=latent_attention_mask)
            step.fill_(t)
            ops.topk_sample_update(logits.squeeze(1), step, rollouts, log_probs, finished, self.decoder.eos_idx, top_k,
                                   temperature, seed=seed_t, uniforms=uniforms)
            if t % self.EOS_POLL_INTERVAL == 0 and bool(finished.all()):
                break
        return self._finish_rollouts(rollouts, log_probs)

    forward_rollout_policy = cached_forward_rollout_policy

    def prepare_rollouts_for_policy_theta(self, rollouts, rollout_mask):
        """models.py:949-958: right-shifted rollouts and their True = padding attention mask."""
        lens = rollout_mask.sum(dim=-1, keepdim=True) - 1
        pos = torch.arange(int(torch.max(lens)), device=rollouts.device).repeat([rollouts.shape[0], 1])
        return rollouts[:, :-1], pos >= lens

    def forward_teacher_forced(self, img_latent, latent_attention_mask, lmx_seqs, checkpoint_grads=False):
        """models.py:963-967."""
        input_seqs, target_seqs, lmx_attention_mask = batchify_and_split_lmx_seqs(lmx_seqs, self.decoder.pad_idx, img_latent.device)
        pred = self.decoder(input_seqs, img_latent, lmx_attention_mask, latent_attention_mask, checkpoint_grads=checkpoint_grads)
        return pred, target_seqs

    def batch_policy_inference(self, imgs, max_actions, top_k, temperature):
        """models.py:971-974 (one rollout per image; the caller applies the transition head as the reference does)."""
        img_latent, latent_attention_mask = self.encoder(imgs)
        return self.forward_rollout_policy(img_latent, latent_attention_mask, max_actions, top_k, temperature)


class OMRCELoss(nn.Module):
    """models.py:784-796: token-level cross entropy over non-pad targets (acai_ce_loss forward + backward)."""

    def __init__(self, pad_idx, label_smoothing=0.0):
        super().__init__()
        self.pad_idx, self.label_smoothing = pad_idx, float(label_smoothing)

    def forward(self, pred, target_seqs):
        return F.ce_loss(pred, target_seqs.to(pred.device), self.pad_idx, label_smoothing=self.label_smoothing)[0]


def default_vocab_path():
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data", "lmx_vocab.txt")
