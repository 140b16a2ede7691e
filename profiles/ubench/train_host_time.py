"""Host time of one teacher-forced training step: the full-DEPTH model (12 + 12 layers, so the launch count of the bench's
step) at a tiny width and on tiny inputs, where the kernels are negligible - wall time per step is then what the host
needs to enqueue a step.  Compare with the bench's GPU-bound 183 ms."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from acai_omr_b200 import ops, synth, train  # noqa: E402
from acai_omr_b200.models import models as M  # noqa: E402

dev = "cuda"
enc = M.FineTuneOMREncoder(16, 60, 200, 12, num_layers=12, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.05)
dec = M.OMRDecoder(1536, M.default_vocab_path(), num_layers=12, hidden_dim=128, num_heads=2, mlp_dim=256, transformer_dropout=0.1)
vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=128, transition_head_dropout=0.05).to(dev).train()
groups, _ = vit.create_fine_tune_param_groups(1e-4, 1e-4, 0.9)
tr = train.FlatTrainer(vit, groups)
imgs = [synth.synth_system(32, 64, i).to(dev) for i in range(64)]
batch = list(zip(imgs, [synth.synth_lmx(24, i).to(dev) for i in range(64)]))
for _ in range(3):
    train.teacher_forced_step(vit, batch, tr)
torch.cuda.synchronize()
n0 = ops.launch_count()
t0 = time.perf_counter()
for _ in range(5):
    train.teacher_forced_step(vit, batch, tr)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 5
print(f"{(ops.launch_count() - n0) // 5} launches per step, {dt * 1e3:.1f} ms per step on a model whose kernels are negligible")
