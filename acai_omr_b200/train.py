"""Data-parallel training plumbing for the two training steps of the hot path:
teacher-forced ViTOMR + OMRCELoss (omr_teacher_force_train.py:104-140) and MAE pre-training + MAELoss.

B200-first layout: all trainable parameters of a module live in ONE flat fp32 buffer, their gradients in a
second, AdamW moments in a third and fourth, and a bf16 shadow of the parameters (what the tcgen05 GEMMs read)
in a fifth.  `p.data` / `p.grad` are views into those buffers, so
  * the gradient exchange is a single NCCL all-reduce over NVLink/NVSwitch of one contiguous buffer (no bucketing
    copies; NVLS in-switch reduction applies when NCCL enables it),
  * the optimiser is one fused kernel launch per parameter group (`acai_adamw_step`), which also refreshes the
    bf16 shadow - no per-weight cast kernels after a step.
The ragged batch is sharded across ranks by striding (`shard`), and losses are normalised by the GLOBAL token /
patch count so an N-rank step computes exactly the gradient of the single-process step on the whole batch.

One process per GPU; `torch.distributed` is only used for the two collectives (count, gradients).
"""
from __future__ import annotations

import ctypes as C

import os

import torch
import torch.distributed as dist

from . import _lib, layers, ops
from . import functional as F

ALIGN = 64  # elements: every parameter starts on a 256 B (fp32) / 128 B (bf16) boundary, as TMA wants


def shard(items, rank, world, cost=None):
    """This rank's share of a ragged batch.  Without `cost`: stride sharding (consecutive systems cycle through aspect
    ratios / lengths).  With `cost(item) -> number` (e.g. patch tokens + target tokens): greedy bin packing - items in
    decreasing cost, each to the rank with the smallest running sum (ties to the lowest rank) - so that the per-rank
    sums of sequence lengths are balanced.  Deterministic: every rank computes the same assignment."""
    items = list(items)
    if cost is None or world == 1:
        return items[rank::world]
    costs = [cost(it) for it in items]
    order = sorted(range(len(items)), key=lambda i: (-costs[i], i))
    load, mine = [0] * world, []
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += costs[i]
        if r == rank:
            mine.append(i)
    return [items[i] for i in sorted(mine)]


def _complement(ranges, total):
    """The parts of [0, total) not covered by `ranges` (a list of (start, end))."""
    out, pos = [], 0
    for a, b in sorted(ranges):
        if a > pos:
            out.append((pos, a))
        pos = max(pos, b)
    if pos < total:
        out.append((pos, total))
    return out


def _world(group):
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


class FlatTrainer:
    """Flat parameters + gradients + fused AdamW (+ single-buffer gradient all-reduce).

    `param_groups`: None (all trainable parameters, one group) or the list of {"params": iterable, "lr": float,
    ["weight_decay": float]} dicts that `create_fine_tune_param_groups` (models.py:761-781) returns."""

    def __init__(self, module, param_groups=None, lr=1e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01,
                 process_group=None, shadow_dtype=torch.bfloat16):
        self.module, self.group = module, process_group
        self.betas, self.eps, self.step_count = betas, eps, 0
        if param_groups is None:
            param_groups = [{"params": [p for p in module.parameters() if p.requires_grad], "lr": lr}]
        seen, self.param_groups, self.params, self.offsets = set(), [], [], []
        off = 0
        for g in param_groups:
            plist = [p for p in g["params"] if p.requires_grad and id(p) not in seen]
            seen.update(id(p) for p in plist)
            start = off
            for p in plist:
                self.params.append(p)
                self.offsets.append(off)
                off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
            self.param_groups.append({"lr": float(g.get("lr", lr)), "weight_decay": float(g.get("weight_decay", weight_decay)),
                                      "start": start, "end": off, "params": plist})
        if not self.params:
            raise ValueError("FlatTrainer: no trainable parameters")
        dev = self.params[0].device      # CPU tensors are accepted for the layout / collectives (gloo tests); step() is CUDA-only
        self.numel = off
        self.flat_param = torch.zeros(off, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(off, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                view = self.flat_param[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p.grad = self.flat_grad[o:o + p.numel()].view(p.shape)
                p._acai_grad_owned = True      # backward GEMMs accumulate into .grad directly (functional._owned_grad)
        self._works, self._reduced = [], []      # gradient all-reduces in flight / flat ranges they cover (this step)
        self.overlapped_elements = 0
        self.shadow = None
        if shadow_dtype is not None:
            self.shadow = self.flat_param.to(shadow_dtype)
            for p, o in zip(self.params, self.offsets):
                layers.WEIGHTS.register_shadow(p, self.shadow[o:o + p.numel()].view(p.shape))

    # ---- one optimisation step --------------------------------------------------------------------
    def zero_grad(self):
        self.flat_grad.zero_()
        for p, o in zip(self.params, self.offsets):       # autograd may have replaced a view (it never does when .grad is set)
            if p.grad is None or p.grad.data_ptr() != self.flat_grad.data_ptr() + 4 * o:
                p.grad = self.flat_grad[o:o + p.numel()].view(p.shape)

    def mean_count(self, local_count):
        """Global count / world size as a 1-element fp32 device tensor: the loss normaliser under data parallelism."""
        c = local_count.detach().reshape(1).to(torch.float32)
        w = _world(self.group)
        if w > 1:
            c = c.clone()
            dist.all_reduce(c, group=self.group)
            c = c / w
        return c

    def allreduce_gradients(self):
        """Sum the flat gradient buffer over ranks; the 1/world is folded into the AdamW kernel.  Whatever
        `allreduce_ready` has not already put in flight goes out now (one collective per remaining contiguous range);
        then the step waits for all of them."""
        if _world(self.group) > 1:
            for a, b in _complement(self._reduced, self.numel):
                self._works.append(dist.all_reduce(self.flat_grad[a:b], group=self.group, async_op=True))
            for w in self._works:
                w.wait()
        self.overlapped_elements = sum(b - a for a, b in self._reduced)     # what went out before the backward had finished
        self._works, self._reduced = [], []

    def allreduce_ready(self, params):
        """Overlap: the gradients of `params` are final (their part of the backward has run) - start their all-reduce now,
        asynchronously, while the rest of the backward keeps the SMs busy.  Called from a tensor hook at a point of the
        backward that dominates those parameters (`grad_ready_hook`).  The flat layout makes this one collective per
        contiguous run of parameters (the decoder + transition head of ViTOMR are one run: 2/3 of the 1.22 GB)."""
        if _world(self.group) <= 1:
            return
        want = {id(p) for p in params}
        runs = []
        for p, o in zip(self.params, self.offsets):
            if id(p) in want:
                end = o + (p.numel() + ALIGN - 1) // ALIGN * ALIGN
                if runs and runs[-1][1] == o:
                    runs[-1][1] = end
                else:
                    runs.append([o, end])
        for a, b in runs:
            if any(a < rb and ra < b for ra, rb in self._reduced):
                raise RuntimeError("FlatTrainer.allreduce_ready: range reduced twice in one step")
            self._reduced.append((a, b))
            self._works.append(dist.all_reduce(self.flat_grad[a:b], group=self.group, async_op=True))

    def grad_ready_hook(self, tensor, modules):
        """Register `allreduce_ready(parameters of modules)` to run when the gradient w.r.t. `tensor` has been computed."""
        if _world(self.group) > 1 and tensor.requires_grad and os.environ.get("ACAI_DP_OVERLAP", "1") != "0":
            params = [p for m in modules for p in m.parameters() if p.requires_grad]
            tensor.register_hook(lambda g: (self.allreduce_ready(params), None)[1])

    def step(self):
        if self.flat_param.device.type != "cuda":
            raise ops.AcaiError("FlatTrainer.step: the fused AdamW kernel needs the parameters on a CUDA device (no CPU fallback)")
        self.step_count += 1
        lib = _lib.load()
        scale = 1.0 / _world(self.group)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for g in self.param_groups:
            s, n = g["start"], g["end"] - g["start"]
            if n == 0:
                continue
            sh = C.c_void_p(self.shadow.data_ptr() + s * self.shadow.element_size()) if self.shadow is not None else None
            ops._call("acai_adamw_step", lib.acai_adamw_step, C.c_void_p(self.flat_param.data_ptr() + 4 * s),
                      C.c_void_p(self.flat_grad.data_ptr() + 4 * s), C.c_void_p(self.exp_avg.data_ptr() + 4 * s),
                      C.c_void_p(self.exp_avg_sq.data_ptr() + 4 * s), sh, n, g["lr"], self.betas[0], self.betas[1], self.eps,
                      g["weight_decay"], self.step_count, scale, st, nbytes=30.0 * n)
        layers.WEIGHTS.bump_generation()

    # ---- checkpoint / resume --------------------------------------------------------------------------
    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "lrs": [g["lr"] for g in self.param_groups], "numel": self.numel}

    def load_state_dict(self, sd):
        if sd["numel"] != self.numel:
            raise ValueError("FlatTrainer.load_state_dict: parameter layout mismatch")
        self.step_count = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        for g, lr in zip(self.param_groups, sd["lrs"]):
            g["lr"] = float(lr)


def teacher_forced_step(vit, batch, trainer: FlatTrainer, autocast=True):
    """One data-parallel step of omr_teacher_force_train.py:113-127 on this rank's shard (no gradient accumulation):
    returns the device tensor [sum of token losses, token count] of the shard."""
    trainer.zero_grad()
    # data parallel: once the backward reaches the encoder's output, the decoder's and the transition head's gradients are
    # final - their all-reduce (2/3 of the buffer) runs under the encoder's backward
    vit._acai_latent_hook = lambda lat: trainer.grad_ready_hook(lat, [vit.decoder, vit.transition_head])
    try:
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            pred, tgt = vit(batch)
            tgt = tgt.to(pred.device)
            norm = trainer.mean_count((tgt != vit.decoder.pad_idx).sum())
            loss, sums = F.ce_loss(pred, tgt, vit.decoder.pad_idx, normaliser=norm)
    finally:
        vit._acai_latent_hook = None
    loss.backward()
    trainer.allreduce_gradients()
    trainer.step()
    return sums


def scheduled_sampling_step(vit, batch, trainer: FlatTrainer, teacher_forcing_prob, sample_tau, use_hard_sampling, autocast=True):
    """One data-parallel step of the reference's actual recipe (omr_teacher_force_train.py:113-127 with
    ScheduledSamplingViTOMR.forward_train): two decoder passes, the second on Gumbel-softmax-mixed embeddings."""
    trainer.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        pred, tgt = vit.forward_train(batch, teacher_forcing_prob, sample_tau, use_hard_sampling)
        tgt = tgt.to(pred.device)
        norm = trainer.mean_count((tgt != vit.decoder.pad_idx).sum())
        loss, sums = F.ce_loss(pred, tgt, vit.decoder.pad_idx, normaliser=norm)
    loss.backward()
    trainer.allreduce_gradients()
    trainer.step()
    return sums


def mae_step(mae, batch, trainer: FlatTrainer, autocast=True, noises=None):
    """One data-parallel MAE pre-training step (mae_pre_train.py train loop) on this rank's shard."""
    trainer.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        pred, loss_mask, target = mae(batch, noises=noises)
        norm = trainer.mean_count(loss_mask.sum())
        loss, sums = F.mae_loss(pred, loss_mask, target, normaliser=norm)
    loss.backward()
    trainer.allreduce_gradients()
    trainer.step()
    return sums
