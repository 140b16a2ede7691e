"""Generate `golden_grpo.pt`: outputs of the UNMODIFIED reference (jsnchon/acai-omr @ 5819ae3) for the GRPO pieces of the
hot path, on the model and images already stored in `golden_small.pt`:

  * GRPOViTOMR(...) built from the TeacherForcedViTOMR state dict (exercises convert_teacher_forced_state_dict with a
    partly frozen FineTuneOMREncoder), expand_img_latent_for_rollout, cached_forward_rollout_policy (models.py:840-1049)
    and prepare_rollouts_for_policy_theta;
  * calc_grpo_objective / calc_entropy_bonus (omr_grpo_train.py:240-282) and the gradient of their sum w.r.t. the logits.
    That file cannot be imported here (tensorboard, datasets, olimpic_app ...), so the three function definitions are
    lifted out of its source with `ast` and executed unmodified.

torch.multinomial's random stream cannot be reproduced by another implementation, so for the rollout it is replaced by a
deterministic sampler from the SAME distribution: inverse CDF over the kept tokens in vocabulary order, driven by the
recorded uniforms `u[r, t]` (the patched torch.topk remembers the vocabulary indices of the current step).

Build container only (needs /root/reference):   python tests/golden/make_golden_grpo.py
"""
import ast
import os
import sys
import tempfile

import torch
import torch.nn.functional as F

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
g = G["vitomr"]
c = g["cfg"]
VOCAB = "lmx_vocab.txt"


def build_parts(fine_tune_depth):
    enc = R.FineTuneOMREncoder(c["enc"]["patch_size"], c["enc"]["pe_max_height"], c["enc"]["pe_max_width"], fine_tune_depth,
                               num_layers=c["enc"]["num_layers"], hidden_dim=c["enc"]["hidden_dim"],
                               num_heads=c["enc"]["num_heads"], mlp_dim=c["enc"]["mlp_dim"])
    dec = R.OMRDecoder(c["dec"]["max_lmx_seq_len"], VOCAB, num_layers=c["dec"]["num_layers"], hidden_dim=c["dec"]["hidden_dim"],
                       num_heads=c["dec"]["num_heads"], mlp_dim=c["dec"]["mlp_dim"], use_caching=True, max_batch_size=12,
                       cache_dtype=torch.float)
    head = R.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=c["transition_head_dim"]).transition_head
    return enc, head, dec


# a teacher-forced state dict with ONE frozen and ONE fine-tuned encoder layer, so the conversion has to shift an index
tf_enc, tf_head, tf_dec = build_parts(1)
tf = R.TeacherForcedViTOMR(tf_enc, None, tf_dec, transition_head_dim=c["transition_head_dim"])
src = g["sd"]                                           # golden model: fine_tune_depth 2 = both layers in fine_tune_blocks
tf_sd = {}
for k, v in src.items():
    k2 = k.replace("encoder.fine_tune_blocks.layers.0.", "encoder.frozen_blocks.layers.0.")
    k2 = k2.replace("encoder.fine_tune_blocks.layers.1.", "encoder.fine_tune_blocks.layers.0.")
    tf_sd[k2] = v
tf.load_state_dict(tf_sd)

enc, head, dec = build_parts(1)
policy = R.GRPOViTOMR(enc, head, dec, tf.state_dict()).eval()

group_size, max_actions, top_k, temperature = 3, 14, 5, 1.2
gen = torch.Generator().manual_seed(21)
with torch.no_grad():
    lat, lat_mask = policy.encoder(g["imgs"])
    lat = policy.transition_head(lat)
    xlat, xmask = policy.expand_img_latent_for_rollout(lat, lat_mask, group_size)
R_rows = xlat.shape[0]
uniforms = torch.rand(R_rows, max_actions, generator=gen)

state = {"t": 0, "idx": None}
_topk, _multinomial = torch.topk, torch.multinomial


def topk_recording(x, k, dim=-1, **kw):
    out = _topk(x, k, dim=dim, **kw)
    state["idx"] = out.indices
    return out


def multinomial_inverse_cdf(p, num_samples=1, **kw):
    state["t"] += 1
    idx = state["idx"]
    out = torch.empty(p.shape[0], 1, dtype=torch.long)
    for r in range(p.shape[0]):
        order = sorted(range(p.shape[1]), key=lambda j: int(idx[r, j]))          # kept tokens in vocabulary order
        target, cum, choice = float(uniforms[r, state["t"]]) * float(p[r].sum()), 0.0, order[-1]
        for j in order:
            cum += float(p[r, j])
            if cum > target:
                choice = j
                break
        out[r, 0] = choice
    return out


torch.topk, torch.multinomial = topk_recording, multinomial_inverse_cdf
try:
    with torch.no_grad():
        # bias <eos> until some rollouts stop early and some do not (mask / clip logic on mixed rows)
        base = policy.decoder.unembed.bias[policy.decoder.eos_idx].item()
        for bump in [0.25 * i for i in range(0, 200)]:
            policy.decoder.unembed.bias[policy.decoder.eos_idx] = base + bump
            state["t"] = 0
            rollouts, logp, rmask = policy.cached_forward_rollout_policy(xlat, xmask, max_actions=max_actions, top_k=top_k,
                                                                        temperature=temperature)
            has_eos = (rollouts == policy.decoder.eos_idx).any(dim=1)
            if bool(has_eos.any()) and not bool(has_eos.all()):
                print("eos bias bump", bump, has_eos.tolist())
                break
        else:
            raise SystemExit("no eos bias gives a mixed fixture")
finally:
    torch.topk, torch.multinomial = _topk, _multinomial
shifted, attn_mask = policy.prepare_rollouts_for_policy_theta(rollouts, rmask)

# ---- objective functions lifted unmodified out of omr_grpo_train.py -------------------------------
with open(os.path.join(REF, "acai_omr/train/omr_grpo_train.py")) as f:
    tree = ast.parse(f.read())
want = {"calc_grpo_objective", "calc_policy_theta_entropy", "calc_entropy_bonus"}
ns = {"torch": torch, "F": F}
mod = ast.Module(body=[n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in want], type_ignores=[])
exec(compile(mod, "omr_grpo_train.py", "exec"), ns)
assert want <= set(ns)

theta = R.GRPOViTOMR(*build_parts(1), tf.state_dict()).train()
for m in theta.modules():
    if isinstance(m, torch.nn.Dropout):
        m.p = 0.0
    if isinstance(m, torch.nn.MultiheadAttention):
        m.dropout = 0.0
with torch.no_grad():                                       # a policy that differs from the old one
    for p in theta.decoder.parameters():
        p.add_(0.05 * torch.randn(p.shape, generator=gen))
theta_logits = theta.decoder(shifted, xlat, attn_mask, xmask)
theta_logits.retain_grad()
advantages = torch.randn(R_rows, generator=gen)
epsilon, num_groups = 0.2, R_rows // group_size
obj = ns["calc_grpo_objective"](theta_logits, rollouts, attn_mask, logp, advantages, epsilon, num_groups)
bonus = ns["calc_entropy_bonus"](theta_logits, attn_mask, theta.decoder.vocab_size)
(obj + 0.3 * bonus).backward()

OUT = dict(cfg=dict(group_size=group_size, max_actions=max_actions, top_k=top_k, temperature=temperature, epsilon=epsilon,
                    num_groups=num_groups, entropy_beta=0.3),
           tf_sd={k: v.clone() for k, v in tf.state_dict().items()},
           policy_sd={k: v.detach().clone() for k, v in policy.state_dict().items()},
           theta_sd={k: v.detach().clone() for k, v in theta.state_dict().items()},
           imgs=g["imgs"], latent=lat, latent_mask=lat_mask, xlatent=xlat, xmask=xmask, uniforms=uniforms,
           rollouts=rollouts, log_probs=logp, rollout_mask=rmask, shifted=shifted, attn_mask=attn_mask,
           theta_logits=theta_logits.detach().clone(), advantages=advantages, objective=obj.detach(), bonus=bonus.detach(),
           dlogits=theta_logits.grad.clone(),
           dec_grads={k: p.grad.detach().clone() for k, p in theta.decoder.named_parameters() if p.grad is not None},
           meta=dict(torch=str(torch.__version__), reference="jsnchon/acai-omr@5819ae3"))
path = os.path.join(HERE, "golden_grpo.pt")
torch.save(OUT, path)
print("wrote", path, os.path.getsize(path), "bytes; rollouts", tuple(rollouts.shape), "eos rows", has_eos.tolist())
