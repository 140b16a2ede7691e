"""Synthetic workloads of BASELINE.json's shapes: "pianoform system" images whose
resolutions follow the reference's DynamicResize maths (acai_omr/utils/utils.py:344-354), and
re-randomised weights so that no two layers are equal.  Host-side helpers only (no kernels)."""
from __future__ import annotations

import math

import torch


def dynamic_resize_shape(aspect: int, max_tokens: int, patch: int = 16):
    """(H, W) for an integer aspect ratio a = W // H under a token budget: H = P*floor(sqrt(N/a)), W = a*H."""
    h = patch * int(math.floor(math.sqrt(max_tokens / aspect)))
    return h, aspect * h


def system_shapes(num_systems: int, max_tokens: int = 1024, aspects=(1, 2, 3, 4, 5, 6, 7, 8)):
    """The config-2 mix: the 8 aspect ratios cycled (64 systems -> sum L = 63 664 for N = 1024)."""
    return [dynamic_resize_shape(aspects[i % len(aspects)], max_tokens) for i in range(num_systems)]


def stress_mix_shapes(num_systems: int, budgets=(256, 512, 1024), aspects=(1, 2, 3, 4, 5, 6, 7, 8)):
    """SURVEY 8(d)'s "stress mix": token budgets N in {256, 512, 1024} x the eight aspect ratios (24 distinct shapes,
    L from 192 to 1024) - a heterogeneous batch whose padded form wastes about half of its rows and whose per-rank token
    sums depend on how it is sharded."""
    combos = [(n, a) for n in budgets for a in aspects]
    return [dynamic_resize_shape(combos[(7 * i) % len(combos)][1], combos[(7 * i) % len(combos)][0]) for i in range(num_systems)]


def synth_system(h: int, w: int, seed: int) -> torch.Tensor:
    """White canvas, 2 x 5 staff lines of 2 px, ~60 dark ellipses/stems, Gaussian noise 0.03; fp32 (1,H,W) in [0,1]."""
    g = torch.Generator().manual_seed(1234 + seed)
    img = torch.ones(h, w)
    for staff in (0.25, 0.65):
        for line in range(5):
            y = int(h * (staff + 0.025 * line))
            img[y:y + 2, :] = 0.0
    ys = torch.randint(2, max(3, h - 6), (60,), generator=g)
    xs = torch.randint(2, max(3, w - 6), (60,), generator=g)
    for y, x in zip(ys.tolist(), xs.tolist()):
        img[y:y + 4, x:x + 5] = 0.05
        img[max(0, y - 12):y, x + 4:x + 5] = 0.05
    img = img + 0.03 * torch.randn(h, w, generator=g)
    return img.clamp_(0.0, 1.0).unsqueeze(0).contiguous()


def rerandomise_(module: torch.nn.Module, seed: int = 1):
    """Every parameter re-drawn (weights N(0,0.02), biases N(0,0.01), LayerNorm weight 1+N(0,0.02)) so layers differ."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            if "norm" in name and name.endswith("weight"):
                v = 1.0 + 0.02 * torch.randn(p.shape, generator=g)
            elif name.endswith("bias"):
                v = 0.01 * torch.randn(p.shape, generator=g)
            else:
                v = 0.02 * torch.randn(p.shape, generator=g)
            p.copy_(v.to(p.device, p.dtype))
    return module


def encoder_flops(lens, layers=12, d=768, mlp=3072, patch_feats=256):
    """Algorithmic FLOPs of the encoder forward: per token 2*(3d^2 + d^2 + 2*d*mlp) per layer
    + 2*patch_feats*d for the embedding; attention 4*L^2*d per layer per image (QK^T and PV, 2 FLOP/MAC)."""
    per_tok_layer = 2.0 * (4 * d * d + 2 * d * mlp)
    lin = sum(lens) * (layers * per_tok_layer + 2.0 * patch_feats * d)
    attn = layers * 4.0 * d * sum(l * l for l in lens)
    return lin + attn


def lmx_lengths(num_systems: int, cycle=(384, 512, 640, 768, 896, 1024, 448, 576)):
    """Target LMX sequence lengths (tokens incl. <bos>/<eos>) for the teacher-forced workload: a fixed cycle with mean
    656, inside the reference's MAX_LMX_SEQ_LEN = 1536 (omr_teacher_force_train.py:25)."""
    return [cycle[i % len(cycle)] for i in range(num_systems)]


def synth_lmx(length: int, seed: int, vocab: int = 227, bos: int = 0, eos: int = 2, first: int = 3) -> torch.Tensor:
    g = torch.Generator().manual_seed(4321 + seed)
    body = torch.randint(first, vocab, (length - 2,), generator=g)
    return torch.cat([torch.tensor([bos]), body, torch.tensor([eos])])


def decoder_flops(tgt_lens, mem_lens, layers=12, d=1024, mlp=4096, vocab=227):
    """Algorithmic forward FLOPs of the teacher-forced LMX decoder: per target token and layer 2*(4d^2 self + 2d^2 cross q/o
    + 2*d*mlp); per memory token and layer 2*(2d^2) for the cross K/V projection; causal self-attention 2*T^2*d and
    cross-attention 4*T*S*d per layer per system; vocabulary projection 2*d*vocab per token."""
    T, S = sum(tgt_lens), sum(mem_lens)
    lin = layers * (T * 2.0 * (6 * d * d + 2 * d * mlp) + S * 2.0 * 2 * d * d) + T * 2.0 * d * vocab
    attn = layers * sum(2.0 * t * t * d + 4.0 * t * s * d for t, s in zip(tgt_lens, mem_lens))
    return lin + attn


def peaked_unembed_(decoder, c: float = 0.05, seed: int = 7, emb_scale: float = 2.5):
    """Give a random-init OMRDecoder the peaked next-token distributions of a trained model: the vocabulary projection is
    tied to a permutation of the token-embedding table (unembed.weight[j] = c * emb[perm[j]] / std(emb), the table itself
    scaled by `emb_scale` first), so that the residual stream's memory of the input token produces one dominant logit
    (mean top-1 / top-2 margin about 4.9 at max|logit| about 13 on the full-size decoder; 0.6 % of the steps still have a
    margin under 0.05) while the attention and MLP blocks still move it: different encoder memories give different
    sequences.  `emb_scale` sets how often the memory decides (1.0: every few steps, 3.0: never).  Random-init logits are
    nearly flat (margin 0.2), where greedy decoding is a coin flip for ANY two implementations that differ by one
    rounding; this is the workload on which "reproduces the reference's greedy sequences" can be measured."""
    g = torch.Generator().manual_seed(seed)
    emb = decoder.vocab_embedding.weight
    with torch.no_grad():
        emb.mul_(emb_scale)
        perm = torch.randperm(emb.shape[0], generator=g).to(emb.device)
        decoder.unembed.weight.copy_(c * emb[perm] / emb.std().clamp_min(1e-12))
    return decoder
