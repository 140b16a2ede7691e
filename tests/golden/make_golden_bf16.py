"""Generate `golden_bf16.pt`: the UNMODIFIED reference (jsnchon/acai-omr @ 5819ae3) in its real inference mode.

    python tests/golden/make_golden_bf16.py          (build container only: needs /root/reference)

`acai_omr/inference/vitomr_inference.py:79-86,94` runs the encoder in fp32 OUTSIDE autocast, then the transition head
and `ViTOMR.cached_greedy_generate` under `autocast(dtype=torch.bfloat16)` with bf16 KV caches
(`decoder.to_cached_version(MAX_BATCH_SIZE, torch.bfloat16)`).  This script runs exactly that call sequence on CPU
(`autocast(device_type="cpu")`, torch 2.11) on a head_dim-64 model and records the per-step logits the greedy loop saw
(by wrapping `decoder.cached_generate`, nothing in the reference is edited), the token sequences, the log-probs and the
head output.  The oracle's bf16 mode (`oracle.vitomr_oracle.autocast_bf16`) is pinned to these; the CUDA-graph decode
engine (head_dim 64 -> the TMA / split-K kernels bench.py times) is compared with them directly.
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
os.chdir(REF)

from acai_omr.models import models as R  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_bf16.pt")
VOCAB = "lmx_vocab.txt"


def rerandomise(module, seed, peaked=1.0):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            if ".norm" in name and name.endswith("weight") or name.endswith("norm.weight"):
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            elif name.endswith("bias"):
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
            elif "pos_embedding" in name:
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif name.startswith("decoder.unembed") or name.startswith("unembed"):
                p.copy_(peaked * 0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
    return module


def sd_of(m):
    """transition head + decoder weights only (the encoder does not take part: `latent` stands for its output)."""
    return {k: v.detach().clone() for k, v in m.state_dict().items() if not k.startswith("encoder.")}


def run(name, peaked, seed, B=6, S=20, max_len=24, layers=3, D=128, H=2, mlp=256, enc_dim=96, head_dim=192):
    torch.manual_seed(seed)
    enc = R.FineTuneOMREncoder(16, 4, 8, 1, num_layers=1, hidden_dim=enc_dim, num_heads=2, mlp_dim=128)
    dec = R.OMRDecoder(max_len + 8, VOCAB, num_layers=layers, hidden_dim=D, num_heads=H, mlp_dim=mlp, use_caching=True,
                       max_batch_size=B, cache_dtype=torch.bfloat16)
    vit = rerandomise(R.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=head_dim), seed + 1, peaked).eval()
    g = torch.Generator().manual_seed(seed + 2)
    lens = [S, S - 3, 7, S, 12, 1][:B]
    latent = torch.randn(B, S, enc_dim, generator=g)                      # stands for the fp32 encoder output
    mask = torch.arange(S)[None, :] >= torch.tensor(lens)[:, None]
    step_logits = []
    inner = vit.decoder.cached_generate

    def recording(token_t, time_step, latent_attention_mask=None):
        out = inner(token_t, time_step, latent_attention_mask=latent_attention_mask)
        step_logits.append(out.detach().clone().squeeze(1))
        return out
    vit.decoder.cached_generate = recording
    with torch.no_grad(), torch.autocast(device_type="cpu", dtype=torch.bfloat16):
        head = vit.transition_head(latent)
        seqs, log_probs, seq_mask = vit.cached_greedy_generate(head, mask, max_len=max_len)
    logits = torch.stack(step_logits, dim=1)                              # (B, steps, V), bf16
    print(name, "head", head.dtype, "logits", logits.dtype, tuple(logits.shape), "seqs", tuple(seqs.shape),
          "k_cache", vit.decoder.decoder_blocks.self_attn_caches[0].k_cache.dtype)
    return dict(cfg=dict(max_lmx_seq_len=max_len + 8, num_layers=layers, hidden_dim=D, num_heads=H, mlp_dim=mlp,
                         enc_dim=enc_dim, head_dim=head_dim, max_len=max_len, batch=B),
                sd=sd_of(vit), latent=latent, mask=mask, lens=lens, head=head, logits=logits, seqs=seqs,
                log_probs=log_probs, seq_mask=seq_mask)


if __name__ == "__main__":
    G = {"flat": run("flat", 1.0, 11), "peaked": run("peaked", 20.0, 21)}
    torch.save(G, OUT)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")
