import sys, time, os
sys.path.insert(0, os.getcwd())
import torch
import bench
from acai_omr_b200.pipeline import EncoderPipeline
dev = torch.device("cuda:0")
enc = bench.build_encoder(dev)
imgs = [t.pin_memory() for t in bench.host_images(bench.NUM_SYSTEMS, seed0=0)]
def sync_step():
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        d = [t.to(dev, non_blocking=True) for t in imgs]
        out, mask = enc(d)
    return out, mask
for _ in range(3): sync_step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5): sync_step()
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print("no-D2H loop: host enqueue %.1f ms/step, total %.1f ms/step" % ((t1 - t0) * 200, (t2 - t0) * 200))
pipe = EncoderPipeline(enc, dev)
for _ in pipe.run([imgs] * 3): pass
torch.cuda.synchronize()
ts = [time.perf_counter()]
for _ in pipe.run(imgs for _ in range(8)):
    ts.append(time.perf_counter())
torch.cuda.synchronize(); ts.append(time.perf_counter())
print("pipeline yield times (ms):", [round((b - a) * 1e3, 1) for a, b in zip(ts, ts[1:])])
print("mem reserved GiB", torch.cuda.memory_reserved() / 2**30, "num alloc retries", torch.cuda.memory_stats()["num_alloc_retries"], "segments", torch.cuda.memory_stats()["segment.all.current"])
# variant: D2H on the compute stream (no side stream), H2D on side stream
h = torch.empty(64 * 1024 * 768 * 2, dtype=torch.float32).pin_memory()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5):
    out, mask = sync_step()
    h[:out.numel()].view(out.shape).copy_(out, non_blocking=True)
torch.cuda.synchronize(); print("sync-stream D2H loop %.1f ms/step" % ((time.perf_counter() - t0) * 200))
s2 = torch.cuda.Stream()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5):
    out, mask = sync_step()
    ev = torch.cuda.Event(); ev.record()
    with torch.cuda.stream(s2):
        s2.wait_event(ev)
        h[:out.numel()].view(out.shape).copy_(out, non_blocking=True)
        out.record_stream(s2)
torch.cuda.synchronize(); print("side-stream D2H loop %.1f ms/step" % ((time.perf_counter() - t0) * 200))
