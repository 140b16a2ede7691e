"""Whole encoder layers (12, full size, 64 systems) queued back to back, one pair of events around them: the time that
matters, launch gaps included.  ACAI_FUSE_LINEAR_LN / ACAI_STREAM / ACAI_PDL select the variant."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import layers, ops, synth  # noqa: E402
from acai_omr_b200.models import models as M  # noqa: E402

dev = "cuda"
torch.manual_seed(0)
enc = M.FineTuneOMREncoder(16, 60, 200, 12, num_layers=12, hidden_dim=768, num_heads=12, mlp_dim=3072)
synth.rerandomise_(enc, 1)
enc = enc.to(dev).eval()
imgs = [synth.synth_system(h, w, i).to(dev) for i, (h, w) in enumerate(synth.system_shapes(64))]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def step():
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        return enc.encode_packed(imgs)[0]


for _ in range(4):
    step()
torch.cuda.synchronize()
ts = []
for _ in range(12):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); step(); e1.record()
    ts.append((e0, e1))
torch.cuda.synchronize()
ms = sorted(a.elapsed_time(b) for a, b in ts)
with ops.Profile() as prof:
    step()
k = prof.summary()
tot = sum(v["ms"] for v in k.values())
print(f"fuse={os.environ.get('ACAI_FUSE_LINEAR_LN', '1')} stream={os.environ.get('ACAI_STREAM', 'fp32')} pdl={os.environ.get('ACAI_PDL', '1')}: "
      f"median step {ms[len(ms) // 2]:.3f} ms (min {ms[0]:.3f}); sum of kernels {tot:.3f} ms")
