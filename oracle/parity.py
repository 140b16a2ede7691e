"""Parity metrics shared by tests/ and bench.py's checker legs.  TEST INFRASTRUCTURE (see vitomr_oracle.py's header).

North-star tolerances (BASELINE.json): logits within 1e-3 relative in fp32, 2e-2 in bf16; greedy tokens bit-exact in
fp32, and in bf16 >= 90 % of the reference sequences - where a sequence that departs from the reference only at a step
whose reference top-1 / top-2 logit margin is inside the bf16 tolerance band is a tie flip, not an error (SURVEY 7:
"first-divergence step + logit margin").  Every function takes CPU or CUDA tensors and returns plain Python numbers.
"""
from __future__ import annotations

import torch

FP32_TOL = 1e-3
BF16_TOL = 2e-2
TIE_ULPS = 2.0      # see divergence_report


def bf16_ulp(v):
    """Spacing of bf16 numbers at magnitude v (8 significant bits): 2 ** (floor(log2 v) - 7).  The reference's logits ARE
    bf16 numbers in its inference mode (the unembed Linear runs under autocast), so two candidates whose logits differ by
    at most this much are one rounding apart: which of them argmax returns is decided by the last bit of one
    implementation's accumulation order, not by the algorithm."""
    v = torch.as_tensor(v, dtype=torch.float32).abs().clamp_min(1e-30)
    return torch.exp2(torch.floor(torch.log2(v)) - 7.0)


def errors(x, ref, rows=None):
    """Relative errors of x against ref (optionally on the boolean row selection `rows` over the leading dims):
    max_rel  = max|x - ref| / max|ref|                       (the max-norm figure the round-1 tests gated)
    l2_rel   = ||x - ref|| / ||ref||
    elem_rel = max over elements of |x - ref| / (|ref| + rms(ref))   (per element; rms(ref) keeps zeros of ref finite)
    """
    x, ref = x.detach().float().cpu(), ref.detach().float().cpu()
    if rows is not None:
        rows = rows.cpu()
        x, ref = x[rows], ref[rows]
    d = (x - ref).abs()
    rms = ref.pow(2).mean().sqrt().clamp_min(1e-12)
    return {"max_rel": (d.max() / ref.abs().max().clamp_min(1e-12)).item(),
            "l2_rel": (d.pow(2).sum().sqrt() / ref.pow(2).sum().sqrt().clamp_min(1e-12)).item(),
            "elem_rel": (d / (ref.abs() + rms)).max().item(),
            "max_abs": d.max().item(), "ref_max": ref.abs().max().item(), "n": int(ref.numel())}


def argmax_agreement(logits, ref_logits, tol, valid=None):
    """Step-wise (teacher-forced) token agreement between two (B, T, V) logit tensors computed on the SAME prefixes.
    A disagreement at a step where the reference's top-1 / top-2 margin is <= 2 * tol * max|ref row| is a tie flip: both
    sides may be `tol * max|row|` away from the exact logits.  -> dict(agree, agree_or_tie, worst_margin_over_band)."""
    a, r = logits.detach().float().cpu(), ref_logits.detach().float().cpu()
    same = a.argmax(-1) == r.argmax(-1)
    top2 = r.topk(2, dim=-1).values
    margin = top2[..., 0] - top2[..., 1]
    band = 2.0 * tol * r.abs().amax(dim=-1)
    tie = margin <= band
    if valid is None:
        valid = torch.ones_like(same)
    valid = valid.cpu().bool()
    n = max(int(valid.sum()), 1)
    bad = valid & ~same & ~tie
    ratio = (margin / band.clamp_min(1e-12))[valid & ~same]
    return {"agree": float((same & valid).sum()) / n, "agree_or_tie": float(((same | tie) & valid).sum()) / n,
            "not_tie_disagreements": int(bad.sum()), "steps": n,
            "worst_margin_over_band": float(ratio.max()) if ratio.numel() else 0.0}


def divergence_report(seqs, ref_seqs, ref_step_logits, tol, pad_idx=1):
    """Free-running greedy sequences against the reference's: per-sequence exact match rate, and for every sequence that
    departs, the first step t where it does and the reference's top-1 / top-2 margin at that step against (a) the
    tolerance band 2 * tol * max|logit row| (`tie_flip`) and (b) the resolution of bf16 at the row's largest logit
    (`within_bf16_resolution`: margin <= TIE_ULPS = 2 bf16 ulps.  Two faithful bf16 implementations each deliver a logit
    within one rounding step of the other's, so a pair of candidates two steps apart can swap; the unmodified reference
    under CPU autocast and this repo's CPU restatement of the same semantics already part at margins of 0, 1 and 2 ulps,
    tests/test_oracle.py::test_oracle_bf16_mode_matches_reference_recorded_run).
    `ref_step_logits[:, t - 1]` are the reference logits that chose token t."""
    seqs, ref_seqs = seqs.cpu(), ref_seqs.cpu()
    B = ref_seqs.shape[0]
    T = max(seqs.shape[1], ref_seqs.shape[1])

    def pad(s):
        out = torch.full((B, T), pad_idx, dtype=torch.long)
        out[:, : s.shape[1]] = s
        return out
    a, r = pad(seqs), pad(ref_seqs)
    lg = ref_step_logits.detach().float().cpu()
    rows, exact, ties, unresolved = [], 0, 0, 0
    for b in range(B):
        diff = (a[b] != r[b]).nonzero()
        if diff.numel() == 0:
            exact += 1
            continue
        t = int(diff[0])
        row = lg[b, min(max(t - 1, 0), lg.shape[1] - 1)]
        top = row.topk(2).values
        margin, band = float(top[0] - top[1]), 2.0 * tol * float(row.abs().max())
        ulp = float(bf16_ulp(row.abs().max()))
        ties += margin <= band
        unresolved += margin <= TIE_ULPS * ulp
        rows.append({"seq": b, "first_divergence_step": t, "ref_margin": margin, "tie_band": band,
                     "tie_flip": margin <= band, "bf16_ulp": ulp, "within_bf16_resolution": margin <= TIE_ULPS * ulp})
    tok = float((a == r).float().mean())
    return {"sequences": B, "exact": exact, "exact_rate": exact / B, "exact_or_tie_flip_rate": (exact + ties) / B,
            "exact_or_within_bf16_resolution_rate": (exact + unresolved) / B,
            "token_match_rate": tok, "divergences": rows}
