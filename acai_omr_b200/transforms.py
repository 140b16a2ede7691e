"""Input pipeline in front of the encoder: the reference's `DynamicResize` (acai_omr/utils/utils.py:334-370) preceded by
`v2.ToImage()` + `v2.ToDtype(torch.float32, scale=True)` (omr_teacher_force_train.py:278-282, pre_train.py:162-166), on
the GPU and straight from uint8 scans.

`DynamicResize` keeps the reference's constructor and `forward(img)` (a (C, H, W) tensor in, the resized (C, h, w) fp32
tensor out, clamped to [0, 1]); `resize_batch` does a whole ragged batch in two kernel launches (acai_dynamic_resize) and
returns the list of fp32 CUDA images `Encoder.forward` takes.  The target size and the centre crop are pure shape logic
and stay in Python, exactly as the reference computes them; the arithmetic is the library's (no CPU fallback)."""
from __future__ import annotations

import math
import os

import torch
from torch import nn

from . import _lib, ops
from ._lib import AcaiError, ResizeDesc

RESIZE_F32, RESIZE_U8 = 0, 1


def target_size(height, width, patch_size, max_seq_len):
    """(target_height, target_width) of DynamicResize.forward (utils/utils.py:346-354): integer aspect ratio, the short
    side = patch_size * floor(sqrt(max_seq_len / aspect))."""
    if width > height:
        aspect = width // height
        th = patch_size * math.floor(math.sqrt(max_seq_len / aspect))
        return th, th * aspect
    aspect = height // width
    tw = patch_size * math.floor(math.sqrt(max_seq_len / aspect))
    return tw * aspect, tw


def crop_window(th, tw, patch_size, pe_max_height, pe_max_width, crop_imgs):
    """(top, left, out_h, out_w): the centre crops of utils/utils.py:363-368 (torchvision center_crop rounds the offset with
    Python's round, i.e. half to even)."""
    top, left, oh, ow = 0, 0, th, tw
    if crop_imgs:
        if th / patch_size > pe_max_height:
            oh = pe_max_height * patch_size
            top = int(round((th - oh) / 2.0))
        if tw / patch_size > pe_max_width:
            ow = pe_max_width * patch_size
            left = int(round((tw - ow) / 2.0))
    return top, left, oh, ow


def resize_batch(images, patch_size, max_seq_len, pe_max_height, pe_max_width, crop_imgs):
    """Ragged list of (1, H, W) / (H, W) CUDA tensors, uint8 (scaled by 1/255) or fp32 -> list of (1, h, w) fp32 CUDA tensors."""
    if len(images) == 0:
        return []
    dev = images[0].device
    if dev.type != "cuda":
        raise AcaiError("resize_batch: images must be CUDA tensors (there is no CPU fallback)")
    plans, keep = [], []
    tmp_total = out_total = 0
    for t in images:
        if t.dim() == 3:
            if t.shape[0] != 1:
                raise AcaiError("resize_batch: single-channel images only (NUM_CHANNELS = 1, models.py:11)")
            t = t[0]
        if t.dim() != 2 or t.device != dev:
            raise AcaiError("resize_batch: expected (1, H, W) or (H, W) tensors on one device")
        if t.dtype not in (torch.uint8, torch.float32):
            raise AcaiError(f"resize_batch: uint8 or float32 images, got {t.dtype}")
        if t.stride(-1) != 1:
            t = t.contiguous()
        H, W = t.shape
        th, tw = target_size(H, W, patch_size, max_seq_len)
        if th <= 0 or tw <= 0:
            raise ValueError(f"image of {H} x {W} pixels has no valid target size for max_seq_len {max_seq_len}")
        top, left, oh, ow = crop_window(th, tw, patch_size, pe_max_height, pe_max_width, crop_imgs)
        plans.append((t.data_ptr(), RESIZE_U8 if t.dtype == torch.uint8 else RESIZE_F32, H, W, t.stride(0), tmp_total, out_total,
                      th, tw, top, left, oh, ow))
        tmp_total += H * ow
        out_total += (oh * ow + 3) & ~3                      # every image starts on a 16-byte boundary
        keep.append(t)
    # ONE scratch and ONE output allocation for the whole batch (the returned images are views), one descriptor upload
    tmp = torch.empty(max(tmp_total, 1), dtype=torch.float32, device=dev)
    flat = torch.empty(out_total, dtype=torch.float32, device=dev)
    tmp_base, out_base = tmp.data_ptr(), flat.data_ptr()
    arr = (ResizeDesc * len(images))()
    outs = []
    for i, (src, sdt, H, W, stride, toff, ooff, th, tw, top, left, oh, ow) in enumerate(plans):
        arr[i] = ResizeDesc(src, sdt, H, W, stride, tmp_base + 4 * toff, out_base + 4 * ooff, th, tw, top, left, oh, ow, ow)
        outs.append(flat[ooff:ooff + oh * ow].view(1, oh, ow))
    host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
    descs = ops.upload_small(host, dev)
    # widest filter window of the batch (either axis): support = 2 * max(in / out, 1) on each side of the centre
    taps = max(2 * math.ceil(2.0 * max(p[2] / p[7], p[3] / p[8], 1.0)) + 2 for p in plans)
    use_table = os.environ.get("ACAI_RESIZE_TABLE", "1") != "0"
    ops._call("acai_dynamic_resize", _lib.load().acai_dynamic_resize, ops._p(descs), len(images),
              max(p[2] for p in plans), max(p[11] for p in plans), max(p[12] for p in plans), taps if use_table else 0, ops._stream())
    ops._bump()                       # two kernels: rows pass, columns pass
    return outs


class DynamicResize(nn.Module):
    """acai_omr.utils.utils.DynamicResize (utils/utils.py:334-370): same constructor, same forward; also accepts the uint8
    image `v2.ToImage()` produces (the 1/255 scaling of `v2.ToDtype(float32, scale=True)` is then applied on load)."""

    def __init__(self, patch_size, max_seq_len, pe_max_height, pe_max_width, crop_imgs):
        super().__init__()
        self.patch_size = patch_size
        self.max_seq_len = max_seq_len
        self.pe_max_height = pe_max_height
        self.pe_max_width = pe_max_width
        self.crop_imgs = crop_imgs

    def forward(self, img):
        return self.forward_batch([img])[0]

    def forward_batch(self, imgs):
        return resize_batch(list(imgs), self.patch_size, self.max_seq_len, self.pe_max_height, self.pe_max_width, self.crop_imgs)
