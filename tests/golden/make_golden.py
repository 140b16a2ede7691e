"""Generate `golden_small.pt` from the UNMODIFIED reference (jsnchon/acai-omr @ 5819ae3).

Run in the build container only (needs /root/reference; the GPU box does not have it):

    python tests/golden/make_golden.py

The reference is imported read-only from /root/reference behind a stub `flask`
package (acai_omr/__init__.py:1 imports Flask, which is not installed).
Everything is CPU fp32, torch 2.11.  Each entry stores the inputs, the weights
(`state_dict`, reference key names) and the reference outputs, so the oracle and
the CUDA modules can be checked without the reference being present.
"""
import os
import sys
import tempfile

import torch

REF = "/root/reference"
stub = tempfile.mkdtemp()
os.makedirs(os.path.join(stub, "flask"))
with open(os.path.join(stub, "flask", "__init__.py"), "w") as f:
    f.write("class Flask: pass\n")
sys.path[:0] = [stub, REF]
sys.dont_write_bytecode = True
os.chdir(REF)  # "lmx_vocab.txt" is a relative path in the reference

from acai_omr.models import models as R  # noqa: E402
from acai_omr.models import kv_caching as RK  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_small.pt")
VOCAB = "lmx_vocab.txt"


def rerandomise(module, seed):
    """nn.TransformerEncoder/Decoder deep-copy their prototype layer, so all layers start identical.
    Re-draw every parameter so a layer-index bug cannot hide."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            if name.endswith("norm.weight") or ".norm" in name and name.endswith("weight"):
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            elif name.endswith("bias"):
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
            elif "pos_embedding" in name or "mask_token" in name:
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(0.15 * torch.randn(p.shape, generator=g))
    return module


def sd_of(m):
    return {k: v.detach().clone() for k, v in m.state_dict().items()}


def imgs_of(shapes, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.rand(1, h, w, generator=g) for h, w in shapes]


G = {}

# ---- 1. ragged encoder (Encoder, eval fast path and train slow path) -------------------------
torch.manual_seed(0)
enc_kwargs = dict(num_layers=3, hidden_dim=32, num_heads=4, mlp_dim=64)
enc = rerandomise(R.Encoder(4, 6, 10, **enc_kwargs), 1)
shapes = [(8, 12), (16, 24), (4, 40), (24, 8)]  # L = 6, 24, 10, 12
imgs = imgs_of(shapes, 2)
enc.eval()
with torch.no_grad():
    emb, mask0 = enc.batchify(imgs)
    y_eval, mask = enc(imgs)  # nested-tensor fast path: pad rows = final-norm bias
    y_single = enc.generate(imgs[1])
enc.train()  # dropout is 0.0 for Encoder -> deterministic slow path
with torch.no_grad():
    y_train, _ = enc(imgs)
G["encoder"] = dict(cfg=dict(patch_size=4, pe_max_height=6, pe_max_width=10, **enc_kwargs), shapes=shapes, seed=2,
                    imgs=imgs, sd=sd_of(enc), batchify=emb, mask=mask, out_eval=y_eval, out_train=y_train,
                    out_single=y_single)

# ---- 2. FineTuneOMREncoder with frozen + fine-tune stacks and interpolated PE ---------------------
ft_kwargs = dict(num_layers=3, hidden_dim=32, num_heads=2, mlp_dim=48, transformer_dropout=0.0)
ft = rerandomise(R.FineTuneOMREncoder(4, 4, 6, 2, **ft_kwargs), 3)
shapes2 = [(8, 16), (8, 40), (24, 12)]  # 2x4 ; 2x10 (w_p > 6 -> interpolated) ; 6x3 (h_p > 4 -> interpolated)
imgs2 = imgs_of(shapes2, 4)
ft.eval()
with torch.no_grad():
    emb2, _ = ft.batchify(imgs2)
    y2, mask2 = ft(imgs2)
    pe_interp = ft.interpolate_pe(6, 3)
G["finetune_encoder"] = dict(cfg=dict(patch_size=4, pe_max_height=4, pe_max_width=6, fine_tune_depth=2, **ft_kwargs),
                             shapes=shapes2, imgs=imgs2, sd=sd_of(ft), batchify=emb2, mask=mask2, out_eval=y2,
                             pe_interp_6x3=pe_interp)

# ---- 3. OMRDecoder: teacher-forced forward, uncached generate, cached steps ---------------------
dec_kwargs = dict(num_layers=3, hidden_dim=32, num_heads=4, mlp_dim=64)
torch.manual_seed(5)
dec_u = rerandomise(R.OMRDecoder(24, VOCAB, **dec_kwargs), 6).eval()
dec_c = R.OMRDecoder(24, VOCAB, **dec_kwargs, use_caching=True, max_batch_size=4, cache_dtype=torch.float).eval()
dec_c.load_state_dict(dec_u.state_dict())
g = torch.Generator().manual_seed(7)
B, S, T = 3, 12, 9
memory = torch.randn(B, S, 32, generator=g)
mem_mask = torch.zeros(B, S, dtype=torch.bool)
mem_mask[0, 9:] = True
mem_mask[2, 5:] = True
tokens = torch.randint(3, dec_u.vocab_size, (B, T), generator=g)
tokens[:, 0] = dec_u.bos_idx
lmx_mask = torch.zeros(B, T, dtype=torch.bool)
lmx_mask[1, 6:] = True
tokens[1, 6:] = dec_u.pad_idx
with torch.no_grad():
    tf_logits = dec_u(tokens, memory, lmx_mask, mem_mask)
    gen_logits = dec_u.generate(tokens, memory, latent_attention_mask=mem_mask)
    dec_c.prepare_caches(memory)
    cached = torch.stack([dec_c.cached_generate(tokens[:, t : t + 1], t, latent_attention_mask=mem_mask).squeeze(1)
                          for t in range(T)], dim=1)
    k_cross0, v_cross0 = dec_c.decoder_blocks.cross_attn_caches[0].get_cached_keys_and_vals()
    k_self2 = dec_c.decoder_blocks.self_attn_caches[2].k_cache[:B, :, :T].clone()
G["decoder"] = dict(cfg=dict(max_lmx_seq_len=24, **dec_kwargs), sd=sd_of(dec_u), memory=memory, mem_mask=mem_mask,
                    tokens=tokens, lmx_mask=lmx_mask, tf_logits=tf_logits, gen_logits=gen_logits, cached_logits=cached,
                    k_cross0=k_cross0.clone(), v_cross0=v_cross0.clone(), k_self2=k_self2)

# ---- 4. ViTOMR end to end: encoder -> head -> cached greedy (fp32), as inference() does ---------
torch.manual_seed(8)
v_enc = R.FineTuneOMREncoder(4, 6, 10, 2, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64)
v_dec = R.OMRDecoder(20, VOCAB, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96,
                     use_caching=True, max_batch_size=4, cache_dtype=torch.float)
vit = rerandomise(R.TeacherForcedViTOMR(v_enc, None, v_dec, transition_head_dim=40), 9).eval()
imgs4 = imgs_of([(8, 12), (16, 24), (12, 40)], 10)
with torch.no_grad():
    lat, lat_mask = vit.encoder(imgs4)
    lat = vit.transition_head(lat)
    # bias the <eos> logit until some rows stop early and some never do, so the fixture exercises
    # mask_and_clip_seqs and the all-finished early exit logic on mixed rows
    base = vit.decoder.unembed.bias[vit.decoder.eos_idx].item()
    for bump in [0.02 * i for i in range(1, 600)]:
        vit.decoder.unembed.bias[vit.decoder.eos_idx] = base + bump
        seqs, logp, smask = vit.cached_greedy_generate(lat, lat_mask, max_len=20)
        has_eos = (seqs == vit.decoder.eos_idx).any(dim=1)
        if bool(has_eos.any()) and not bool(has_eos.all()):
            print("eos bias bump", bump, "rows with eos", has_eos.tolist())
            break
    else:
        raise SystemExit("no eos bias gives a mixed fixture")
    # teacher-forced forward + loss on the same model (dropout off via eval(); encoder on the eval path)
    lmx = [torch.tensor([0, 5, 9, 33, 2]), torch.tensor([0, 7, 2]), torch.tensor([0, 100, 101, 102, 103, 104, 2])]
    pred, tgt = vit(list(zip(imgs4, lmx)))
    loss = R.OMRCELoss(vit.decoder.pad_idx)(pred, tgt)
G["vitomr"] = dict(cfg=dict(enc=dict(patch_size=4, pe_max_height=6, pe_max_width=10, fine_tune_depth=2, num_layers=2,
                                     hidden_dim=32, num_heads=4, mlp_dim=64),
                            dec=dict(max_lmx_seq_len=20, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96),
                            transition_head_dim=40),
                   sd=sd_of(vit), imgs=imgs4, latent=lat, latent_mask=lat_mask, seqs=seqs, log_probs=logp, seq_mask=smask,
                   lmx=lmx, tf_pred=pred, tf_target=tgt, tf_loss=loss)

# ---- 5. MAE forward + loss with recorded noise -----------------------------------------------
torch.manual_seed(11)
mae = rerandomise(R.MAE(0.75, 4, 6, 10, encoder_hidden_dim=32, decoder_hidden_dim=16,
                        encoder_kwargs=dict(num_layers=2, num_heads=4, mlp_dim=64),
                        decoder_kwargs=dict(num_layers=2, num_heads=2, mlp_dim=32)), 12).eval()
imgs5 = imgs_of([(16, 24), (8, 40), (24, 16)], 13)
tgts5 = imgs_of([(16, 24), (8, 40), (24, 16)], 14)
noises = []
_rand = torch.rand


def recording_rand(*a, **k):
    r = _rand(*a, **k)
    noises.append(r.clone())
    return r


torch.rand = recording_rand
try:
    with torch.no_grad():
        pred5, lmask5, target5 = mae(list(zip(imgs5, tgts5)))
finally:
    torch.rand = _rand
loss5 = R.MAELoss()(pred5, lmask5, target5)
G["mae"] = dict(cfg=dict(mask_ratio=0.75, patch_size=4, pe_max_height=6, pe_max_width=10, encoder_hidden_dim=32,
                         decoder_hidden_dim=16, encoder_kwargs=dict(num_layers=2, num_heads=4, mlp_dim=64),
                         decoder_kwargs=dict(num_layers=2, num_heads=2, mlp_dim=32)),
                sd=sd_of(mae), imgs=imgs5, targets=tgts5, noises=noises, pred=pred5, loss_mask=lmask5, target=target5,
                loss=loss5)

# ---- 6. debug_pretrained_mae.pth key list (state_dict layout contract) -------------
dbg = torch.load(os.path.join(REF, "debug_pretrained_mae.pth"), map_location="cpu")
G["debug_mae_keys"] = {k: tuple(v.shape) for k, v in dbg.items()}
G["vitomr_keys"] = {k: tuple(v.shape) for k, v in vit.state_dict().items()}
G["meta"] = dict(torch=torch.__version__, reference="jsnchon/acai-omr@5819ae3")

torch.save(G, OUT)
print("wrote", OUT, os.path.getsize(OUT), "bytes")
