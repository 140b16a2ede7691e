"""Generate `golden_grads.pt`: parameter gradients of the UNMODIFIED reference (jsnchon/acai-omr @ 5819ae3) for the two
training steps of the hot path, on the inputs and weights already stored in `golden_small.pt`:

  * TeacherForcedViTOMR.forward + OMRCELoss (models.py:722-736, 784-796), train mode, every dropout 0;
  * MAE.forward + MAELoss (models.py:249-288) with the recorded torch.rand draws replayed.

Build container only (needs /root/reference):   python tests/golden/make_golden_grads.py
"""
import os
import sys
import tempfile

import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
stub = tempfile.mkdtemp()
os.makedirs(os.path.join(stub, "flask"))
with open(os.path.join(stub, "flask", "__init__.py"), "w") as f:
    f.write("class Flask: pass\n")
sys.path[:0] = [stub, REF]
sys.dont_write_bytecode = True
os.chdir(REF)

from acai_omr.models import models as R  # noqa: E402

G = torch.load(os.path.join(HERE, "golden_small.pt"), map_location="cpu", weights_only=False)
OUT = {}


def grads_of(m):
    return {k: (p.grad.detach().clone() if p.grad is not None else None) for k, p in m.named_parameters()}


# ---- teacher-forced ViTOMR step ---------------------------------------------------------------
g = G["vitomr"]
c = g["cfg"]
enc = R.FineTuneOMREncoder(c["enc"]["patch_size"], c["enc"]["pe_max_height"], c["enc"]["pe_max_width"], c["enc"]["fine_tune_depth"],
                           num_layers=c["enc"]["num_layers"], hidden_dim=c["enc"]["hidden_dim"], num_heads=c["enc"]["num_heads"],
                           mlp_dim=c["enc"]["mlp_dim"], transformer_dropout=0.0)
dec = R.OMRDecoder(c["dec"]["max_lmx_seq_len"], "lmx_vocab.txt", num_layers=c["dec"]["num_layers"], hidden_dim=c["dec"]["hidden_dim"],
                   num_heads=c["dec"]["num_heads"], mlp_dim=c["dec"]["mlp_dim"], transformer_dropout=0.0)
vit = R.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=c["transition_head_dim"], transition_head_dropout=0.0)
vit.load_state_dict(g["sd"])
for p in vit.parameters():          # exercise every gradient, including the ones the fine-tune recipe freezes
    p.requires_grad_(True)
vit.train()
pred, tgt = vit(list(zip(g["imgs"], g["lmx"])))
loss = R.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
loss.backward()
OUT["vitomr"] = dict(loss=loss.detach().clone(), grads=grads_of(vit))

# ---- MAE step ---------------------------------------------------------------------------------
g = G["mae"]
c = g["cfg"]
mae = R.MAE(c["mask_ratio"], c["patch_size"], c["pe_max_height"], c["pe_max_width"], c["encoder_hidden_dim"],
            c["decoder_hidden_dim"], c["encoder_kwargs"], c["decoder_kwargs"])
mae.load_state_dict(g["sd"])
mae.train()
replay = [n.clone() for n in g["noises"]]
_rand = torch.rand
torch.rand = lambda *a, **k: replay.pop(0)
try:
    pred, lmask, target = mae(list(zip(g["imgs"], g["targets"])))
finally:
    torch.rand = _rand
assert torch.equal(lmask, g["loss_mask"])
loss = R.MAELoss()(pred, lmask, target)
loss.backward()
OUT["mae"] = dict(loss=loss.detach().clone(), grads=grads_of(mae))
# ---- scheduled-sampling step (ScheduledSamplingViTOMR.forward_train, models.py:819-834) -----------------------
g = G["vitomr"]
c = g["cfg"]


def build_ss():
    enc = R.FineTuneOMREncoder(c["enc"]["patch_size"], c["enc"]["pe_max_height"], c["enc"]["pe_max_width"], c["enc"]["fine_tune_depth"],
                               num_layers=c["enc"]["num_layers"], hidden_dim=c["enc"]["hidden_dim"], num_heads=c["enc"]["num_heads"],
                               mlp_dim=c["enc"]["mlp_dim"], transformer_dropout=0.0)
    dec = R.OMRDecoder(c["dec"]["max_lmx_seq_len"], "lmx_vocab.txt", num_layers=c["dec"]["num_layers"], hidden_dim=c["dec"]["hidden_dim"],
                       num_heads=c["dec"]["num_heads"], mlp_dim=c["dec"]["mlp_dim"], transformer_dropout=0.0)
    m = R.ScheduledSamplingViTOMR(enc, None, dec, transition_head_dim=c["transition_head_dim"], transition_head_dropout=0.0)
    m.load_state_dict(g["sd"])
    for p in m.parameters():
        p.requires_grad_(True)
    return m.train()


OUT["scheduled_sampling"] = {}
for hard in (False, True):
    ss = build_ss()
    B, T, V = len(g["lmx"]), max(int(s.shape[0]) for s in g["lmx"]) - 1, ss.decoder.vocab_size
    torch.manual_seed(77)
    # the two draws forward_train makes, in its order: torch.rand((B,T)) then empty_like(logits).exponential_() (F.gumbel_softmax)
    rand = torch.rand((B, T))
    gumbel = -torch.empty((B, T, V)).exponential_().log()
    torch.manual_seed(77)
    pred, tgt = ss.forward_train(list(zip(g["imgs"], g["lmx"])), 0.5, 1.5, hard)
    loss = R.OMRCELoss(ss.decoder.pad_idx)(pred, tgt)
    loss.backward()
    OUT["scheduled_sampling"]["hard" if hard else "soft"] = dict(
        tf_prob=0.5, tau=1.5, rand=rand, gumbel=gumbel, pred=pred.detach().clone(), target=tgt.clone(),
        loss=loss.detach().clone(), grads=grads_of(ss))
    print("scheduled sampling", "hard" if hard else "soft", "loss", float(loss), "sampled positions", int((rand < 0.5).sum()))

OUT["meta"] = dict(torch=str(torch.__version__), reference="jsnchon/acai-omr@5819ae3")

path = os.path.join(HERE, "golden_grads.pt")
torch.save(OUT, path)
for k in ("vitomr", "mae"):
    print(k, "loss", float(OUT[k]["loss"]), "params", len(OUT[k]["grads"]), "without grad",
          [n for n, v in OUT[k]["grads"].items() if v is None])
print("wrote", path, os.path.getsize(path), "bytes")
