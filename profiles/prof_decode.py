"""KV-cached greedy decode of the full-size LMX decoder (12 layers, 1024-d, 16 heads, vocab 227), bf16, random-init
weights, synthetic encoder memory of the config-2 length mix.  Prints ms/step and tokens/s per batch size.
    python profiles/prof_decode.py [B ...] [--steps N]
Under ncu use a small --steps (the step is one replayed CUDA graph)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from acai_omr_b200 import decode_engine, synth  # noqa: E402
from acai_omr_b200.layers import Packed  # noqa: E402
from acai_omr_b200.models import models as M  # noqa: E402


def build(B, max_len, dev="cuda"):
    torch.manual_seed(0)
    dec = M.OMRDecoder(1536, M.default_vocab_path(), num_layers=12, hidden_dim=1024, num_heads=16, mlp_dim=4096,
                       use_caching=True, max_batch_size=1, cache_dtype=torch.bfloat16)   # module caches unused by the engine
    synth.rerandomise_(dec, 2)
    dec = dec.to(dev).eval()
    lens = [(h // 16) * (w // 16) for h, w in synth.system_shapes(B)]
    mp = Packed(lens, dev)
    g = torch.Generator(device=dev).manual_seed(3)
    mem = torch.randn(mp.total, 1024, device=dev, generator=g).bfloat16()
    return dec, mem, mp


def run(B, steps):
    max_len = steps + 1
    dec, mem, mp = build(B, max_len)
    holder = type("H", (), {})()
    holder.decoder = dec
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        eng = decode_engine.GreedyDecodeEngine(dec, B, max_len, mp.total, torch.bfloat16, torch.device("cuda"))
        t0 = time.perf_counter()
        eng.prefill(mem, mp)
        torch.cuda.synchronize()
        t_prefill = time.perf_counter() - t0
        eng.generate(stop_early=False)           # captures the graph + one full pass (warm-up)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        seqs, _ = eng.generate(stop_early=False)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"B={B} steps={steps} prefill {t_prefill * 1e3:.1f} ms  decode {ms:.1f} ms  {ms / steps:.3f} ms/step  "
          f"{B * steps / ms * 1e3:.0f} tok/s  first tokens {seqs[0, :6].tolist()}", flush=True)


if __name__ == "__main__":
    args = sys.argv[1:]
    steps = 1023
    if "--steps" in args:
        i = args.index("--steps")
        steps = int(args[i + 1])
        del args[i:i + 2]
    for B in [int(a) for a in args] or [1, 32, 256]:
        run(B, steps)
        torch.cuda.empty_cache()
