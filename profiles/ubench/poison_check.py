"""Debug aid: run the fp32 golden training steps with the caching allocator's free blocks POISONED (filled with NaN bit
patterns), so that any kernel reading memory it assumes initialised produces NaNs / mismatches deterministically."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from acai_omr_b200.models import models as M  # noqa: E402

DEV = "cuda"
golden = torch.load(os.path.join(ROOT, "tests/golden/golden_small.pt"), map_location="cpu", weights_only=False)
gg = torch.load(os.path.join(ROOT, "tests/golden/golden_grads.pt"), map_location="cpu", weights_only=False)


def poison(seed):
    g = torch.Generator().manual_seed(seed)
    blocks = []
    for _ in range(400):
        n = int(torch.randint(1, 1 << 18, (1,), generator=g))
        blocks.append(torch.full((n,), float("nan"), device=DEV))
    for n in (1 << 20, 1 << 22, 1 << 24):
        blocks.append(torch.full((n,), float("nan"), device=DEV))
    torch.cuda.synchronize()
    del blocks


def compare(module, ref, what):
    bad = []
    for name, p in module.named_parameters():
        got = p.grad.detach().float().cpu()
        if not torch.allclose(got, ref[name], atol=3e-5, rtol=2e-3):
            bad.append((name, float((got - ref[name]).abs().max()), bool(torch.isnan(got).any())))
    print(what, "OK" if not bad else bad[:6], flush=True)


def mae_step():
    g, r = golden["mae"], gg["mae"]
    c = g["cfg"]
    mae = M.MAE(c["mask_ratio"], c["patch_size"], c["pe_max_height"], c["pe_max_width"], c["encoder_hidden_dim"],
                c["decoder_hidden_dim"], c["encoder_kwargs"], c["decoder_kwargs"])
    mae.load_state_dict(g["sd"])
    mae = mae.to(DEV).train()
    pred, lmask, target = mae(list(zip([t.to(DEV) for t in g["imgs"]], [t.to(DEV) for t in g["targets"]])), noises=g["noises"])
    loss = M.MAELoss()(pred, lmask, target)
    loss.backward()
    print("mae loss", float(loss), float(r["loss"]))
    compare(mae, r["grads"], "mae grads")


def tf_step():
    g, r = golden["vitomr"], gg["vitomr"]
    c = g["cfg"]
    enc = M.FineTuneOMREncoder(4, 6, 10, 2, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64, transformer_dropout=0.0)
    dec = M.OMRDecoder(20, M.default_vocab_path(), num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96, transformer_dropout=0.0)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=c["transition_head_dim"], transition_head_dropout=0.0)
    vit.load_state_dict(g["sd"])
    for p in vit.parameters():
        p.requires_grad_(True)
    vit = vit.to(DEV).train()
    pred, tgt = vit(list(zip([t.to(DEV) for t in g["imgs"]], [t.to(DEV) for t in g["lmx"]])))
    loss = M.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
    loss.backward()
    print("tf loss", float(loss), float(r["loss"]))
    compare(vit, r["grads"], "tf grads")


for trial in range(4):
    poison(trial)
    mae_step()
    poison(100 + trial)
    tf_step()
