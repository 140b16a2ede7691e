"""Where do the ~1.5 ms between the encoder step's kernel sum and its wall time go?  Times the bench's encoder step
(64 systems, 63 664 tokens, images resident) (a) eagerly, steps queued back to back, (b) as ONE replayed CUDA graph, both
with the L2 flush between steps, and prints the per-kernel sum of a profiling pass.  Diagnostic only: the product path
stays eager (ragged batches rarely repeat a shape signature)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench  # noqa: E402
from acai_omr_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
enc = bench.build_encoder(dev)
imgs = [t.to(dev) for t in bench.host_images(bench.NUM_SYSTEMS, seed0=0)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def step():
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        return enc.encode_packed(imgs)[0]


def timed(fn, n=10):
    ev = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        ev.append((e0, e1))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / n


for _ in range(3):
    step()
torch.cuda.synchronize()
print(f"eager, queued back to back : {timed(step):.3f} ms/step")
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    out = step()
torch.cuda.synchronize()
print(f"one CUDA graph per step    : {timed(g.replay):.3f} ms/step")
print(f"eager again                : {timed(step):.3f} ms/step")
with ops.Profile() as prof:
    step()
s = prof.summary()
print("kernel sum of a profiling pass: %.3f ms" % sum(v["ms"] for v in s.values()))
for k, v in sorted(s.items(), key=lambda kv: -kv[1]["ms"]):
    print(f"  {k:48s} {v['ms']:8.3f} ms  x{v.get('launches', 0)}")
