"""One full-size encoder layer over the config-2 batch (64 ragged systems, 63 664 tokens), twice (warm-up + measured).
Used under ncu:  -k regex:'gemm_bf16|attn_fwd' -s 5 -c 5  profiles the 4 GEMM shapes + the attention of pass 2."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from acai_omr_b200 import synth  # noqa: E402
from acai_omr_b200.models import models as M  # noqa: E402

enc = M.FineTuneOMREncoder(16, 60, 200, 1, num_layers=1)
synth.rerandomise_(enc, 1)
enc = enc.cuda().eval()
imgs = [synth.synth_system(h, w, i).cuda() for i, (h, w) in enumerate(synth.system_shapes(64))]
for _ in range(2):
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        out, pk = enc.encode_packed(imgs)
    torch.cuda.synchronize()
print("ok", out.shape, float(out.float().abs().mean()))
