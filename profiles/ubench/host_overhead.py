"""Where does the HOST time of one encoder step go?  cProfile over 5 resident steps (the GPU work is asynchronous, so
cumulative host time per function = launch overhead, not kernel time)."""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.getcwd())
import torch  # noqa: E402

import bench  # noqa: E402

dev = torch.device("cuda:0")
enc = bench.build_encoder(dev)
imgs = [t.to(dev) for t in bench.host_images(bench.NUM_SYSTEMS, seed0=0)]


def step():
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        return enc.encode_packed(imgs)[0]


for _ in range(3):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host enqueue %.2f ms/step, until idle %.2f ms/step" % ((t1 - t0) * 200, (t2 - t0) * 200))
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
