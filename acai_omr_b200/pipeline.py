"""Host <-> device pipelining around the encoder's public call (Encoder.forward: list of images -> padded fp32 latent
+ mask, models.py:82-139 in the reference).

One synchronous call costs  H2D(images) + compute + D2H(latent): on a B200 the 200 MB padded fp32 latent of a 64-system
batch takes longer to cross PCIe than four encoder layers take to run.  `EncoderPipeline.run` keeps three CUDA streams
busy instead - the images of batch i + 1 are uploaded and the latent of batch i - 1 is downloaded while batch i is in
the kernels - so a serving loop is bounded by max(compute, copies), not their sum.  Results are bit-identical to the
synchronous call (same kernels, same order on the compute stream).
"""
from __future__ import annotations

import collections

import torch


class EncoderPipeline:
    """`depth` batches in flight over `depth + 1` pinned result slots: a yielded (latent, mask) pair stays valid until the
    NEXT result is yielded (the slot a new batch downloads into is the one whose result was handed out two yields ago);
    copy it out if it has to live longer."""

    def __init__(self, encoder, device, depth=2, autocast_dtype=torch.bfloat16):
        if depth < 1:
            raise ValueError("EncoderPipeline: depth must be >= 1")
        self.encoder, self.device, self.depth, self.autocast_dtype = encoder, torch.device(device), depth, autocast_dtype
        self.h2d = torch.cuda.Stream(self.device)
        self.d2h = torch.cuda.Stream(self.device)
        self._slots = [None] * (depth + 1)

    def _host_buffers(self, slot, out, mask):
        cur = self._slots[slot]
        if cur is None or cur[0].numel() < out.numel() or cur[1].numel() < mask.numel() or cur[0].dtype != out.dtype:
            cur = (torch.empty(out.numel(), dtype=out.dtype).pin_memory(), torch.empty(mask.numel(), dtype=mask.dtype).pin_memory())
            self._slots[slot] = cur
        return cur[0][:out.numel()].view(out.shape), cur[1][:mask.numel()].view(mask.shape)

    def _stage(self, batch):
        with torch.cuda.stream(self.h2d):
            dev = [t.to(self.device, non_blocking=True) for t in batch]
            ev = torch.cuda.Event()
            ev.record(self.h2d)
        return dev, ev

    def run(self, batches, before_step=None):
        """Generator over (latent_host, mask_host), one per batch of `batches` (an iterable of lists of pinned CPU image
        tensors), in order.  `before_step()` (optional) is called on the compute stream before each batch's kernels."""
        compute = torch.cuda.current_stream(self.device)
        pending = collections.deque()
        it = iter(batches)
        nxt = next(it, None)
        staged = self._stage(nxt) if nxt is not None else None
        i = 0
        while staged is not None:
            dev, ready = staged
            nxt = next(it, None)
            staged = self._stage(nxt) if nxt is not None else None        # upload of batch i + 1 runs under batch i
            compute.wait_event(ready)
            for t in dev:
                t.record_stream(compute)
            if before_step is not None:
                before_step()
            with torch.no_grad(), torch.autocast("cuda", dtype=self.autocast_dtype, enabled=self.autocast_dtype is not None):
                out, mask = self.encoder(dev)
            done = torch.cuda.Event()
            done.record(compute)
            if len(pending) == self.depth:                                 # frees the slot this batch is about to use
                yield self._finish(pending.popleft())
            out_host, mask_host = self._host_buffers(i % (self.depth + 1), out, mask)
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(done)
                out_host.copy_(out, non_blocking=True)
                mask_host.copy_(mask, non_blocking=True)
                out.record_stream(self.d2h)
                mask.record_stream(self.d2h)
                fin = torch.cuda.Event()
                fin.record(self.d2h)
            pending.append((fin, out_host, mask_host))
            i += 1
        while pending:
            yield self._finish(pending.popleft())

    @staticmethod
    def _finish(entry):
        fin, out_host, mask_host = entry
        fin.synchronize()
        return out_host, mask_host
