"""The GRPO objective of the reference's RL loop (omr_grpo_train.py:239-283) on logits produced by the drop-in decoder.

These are a few row-wise reductions over `theta_logits` (R, T, 227) - a sliver of the step next to the decoder's forward
and backward - written with torch tensor ops so that autograd carries the gradient straight into the CUDA backward of
`acai_omr_b200.functional` (the unembedding GEMM's dX / dW kernels and everything below it).  The rollouts that feed
them come from `GRPOViTOMR.cached_forward_rollout_policy` (device-side top-k sampling, models.py).
Names and argument order follow the reference."""
from __future__ import annotations

import math

import torch


def _chosen_log_probs(theta_logits, tokens):
    """log_softmax(theta_logits)[..., tokens] without materialising the (R, T, V) log-probability tensor twice."""
    z = theta_logits.float()
    return z.gather(-1, tokens.unsqueeze(-1)).squeeze(-1) - torch.logsumexp(z, dim=-1)


def calc_grpo_objective(theta_logits, rollouts, rollout_attention_mask, old_policy_log_probs, advantages, epsilon, num_groups):
    """omr_grpo_train.py:240-262: clipped-ratio objective, averaged per rollout over its valid steps, then over groups.
    theta_logits (R, T-1, V) are the next-token logits of the right-shifted rollouts; rollouts / old_policy_log_probs are
    (R, T); rollout_attention_mask (R, T-1) is True at padding; advantages (R,)."""
    theta_lp = _chosen_log_probs(theta_logits, rollouts[:, 1:])
    ratios = torch.exp(theta_lp - old_policy_log_probs[:, 1:])
    adv = advantages.unsqueeze(1)
    objective = torch.minimum(ratios * adv, ratios.clamp(1 - epsilon, 1 + epsilon) * adv)
    objective = objective.masked_fill(rollout_attention_mask, 0.0)
    lens = (~rollout_attention_mask).sum(dim=-1)
    return (objective.sum(dim=-1) / lens).sum() / num_groups


def calc_policy_theta_entropy(theta_logits, rollout_attention_mask):
    """omr_grpo_train.py:265-275: mean token entropy of every rollout, (R,)."""
    log_p = torch.log_softmax(theta_logits.float(), dim=-1)
    ent = -(log_p.exp() * log_p).sum(dim=-1).masked_fill(rollout_attention_mask, 0.0)
    return ent.sum(dim=-1) / (~rollout_attention_mask).sum(dim=-1)


def calc_entropy_bonus(theta_logits, rollout_attention_mask, vocab_size):
    """omr_grpo_train.py:277-282: mean rollout entropy normalised to [0, 1] by log(vocab)."""
    return calc_policy_theta_entropy(theta_logits, rollout_attention_mask).mean() / math.log(vocab_size)


def group_advantages(raw_group_rewards):
    """omr_grpo_train.py:323-324: group-normalised advantages, flattened to align with the rollouts (R,)."""
    r = raw_group_rewards
    return ((r - r.mean(dim=-1, keepdim=True)) / (r.std(dim=-1, keepdim=True) + 1e-8)).reshape(-1)
