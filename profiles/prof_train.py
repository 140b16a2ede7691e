"""Full-size training steps on one B200 (bf16 autocast, random-init weights, synthetic systems):
    python profiles/prof_train.py tf  [systems]     teacher-forced ViTOMR (12+12 layers) + OMRCELoss + AdamW, reference dropout rates
    python profiles/prof_train.py tf0 [systems]     the same with every dropout 0 (the parity configuration)
    python profiles/prof_train.py mae [systems]     MAE pre-training (12-layer encoder on 25 % + 8-layer decoder) + AdamW
Prints ms/step and a per-entry-point breakdown (CUDA events around every launch of the C-ABI)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from acai_omr_b200 import ops, synth, train  # noqa: E402
from acai_omr_b200.models import models as M  # noqa: E402


def build_tf(dropout=True):
    pe, pd = (0.05, 0.1) if dropout else (0.0, 0.0)         # the reference's rates (omr_teacher_force_train.py:43-46)
    enc = M.FineTuneOMREncoder(16, 60, 200, 12, transformer_dropout=pe)
    dec = M.OMRDecoder(1536, M.default_vocab_path(), num_layers=12, transformer_dropout=pd)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dropout=pe)
    synth.rerandomise_(vit, 1)
    return vit.cuda().train()


def build_mae():
    mae = M.MAE(0.75, 16, 60, 200)
    synth.rerandomise_(mae, 2)
    return mae.cuda().train()


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "tf"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    imgs = [synth.synth_system(h, w, i).cuda() for i, (h, w) in enumerate(synth.system_shapes(n))]
    if kind in ("tf", "tf0"):
        model = build_tf(dropout=kind == "tf")
        batch = list(zip(imgs, [synth.synth_lmx(L, i).cuda() for i, L in enumerate(synth.lmx_lengths(n))]))
        groups, _ = model.create_fine_tune_param_groups(1e-4, 1e-5, 0.9)
        tr = train.FlatTrainer(model, groups)
        step = lambda: train.teacher_forced_step(model, batch, tr)
    else:
        model = build_mae()
        batch = list(zip(imgs, imgs))
        tr = train.FlatTrainer(model, lr=1.5e-4)
        step = lambda: train.mae_step(model, batch, tr)
    print(f"{kind}: {n} systems, {tr.numel / 1e6:.1f} M trainable parameters", flush=True)
    for i in range(2):
        t0 = time.perf_counter()
        sums = step()
        torch.cuda.synchronize()
        print(f"warm-up {i}: {1e3 * (time.perf_counter() - t0):.1f} ms  loss {float(sums[0] / sums[1]):.4f}", flush=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = ops.launch_count()
    e0.record()
    for _ in range(3):
        sums = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"{ms:.2f} ms/step  {n / ms * 1e3:.1f} systems/s  {(ops.launch_count() - l0) // 3} launches/step  "
          f"loss {float(sums[0] / sums[1]):.4f}  peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB", flush=True)
    if os.environ.get("ACAI_PROF_HOST") == "1":           # where the HOST time of a step goes (kernels are asynchronous)
        import cProfile
        import pstats
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            step()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print("host enqueue %.2f ms/step, until idle %.2f ms/step" % ((t1 - t0) * 1e3 / 3, (time.perf_counter() - t0) * 1e3 / 3))
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(2):
            step()
        pr.disable()
        torch.cuda.synchronize()
        pstats.Stats(pr).sort_stats("tottime").print_stats(35)
        return
    with ops.Profile() as prof:
        step()
    rows = sorted(prof.summary().items(), key=lambda kv: -kv[1]["ms"])
    tot = sum(d["ms"] for _, d in rows)
    print(f"per entry point (sum {tot:.2f} ms of kernel time):")
    for name, d in rows[:40]:
        tf = d["flops"] / d["ms"] / 1e9 if d["flops"] else 0.0
        print(f"  {name:58s} x{d['launches']:4d} {d['ms']:9.3f} ms {tf:8.1f} TFLOP/s")


if __name__ == "__main__":
    main()
