"""GPU: parity at BASELINE.json's FULL-SIZE shapes, and of the bf16 CUDA-graph decode engine that bench.py times.

Round-1 parity ran on 32..192-d toys; these tests put the production dimensions (FineTuneOMREncoder(16, 60, 200, 12) =
768 d x 12 layers; OMRDecoder 1024 d x 12 layers x 16 heads, vocab 227 - omr_teacher_force_train.py:23-27,270-275 in the
reference) through the tcgen05 / TMA kernels and compare with the CPU oracle:

  * bf16 (the reference's real inference mode, vitomr_inference.py:79-86,94: head + cached greedy loop under autocast,
    bf16 KV caches): logits within 2e-2 (north star), compared STEP BY STEP on the same token prefix
    (`GreedyDecodeEngine.forced_logits` vs `oracle.cached_forced_logits`), against the oracle's bf16 mode (pinned to the
    unmodified reference's own bf16 run, tests/golden/golden_bf16.pt) and against the fp32 oracle;
  * greedy tokens in bf16: match rate against the oracle, and for every sequence that departs, the first divergence
    step and the oracle's top-1 / top-2 margin there.  GATE (stated): every sequence is reproduced exactly OR departs
    first at a step where the oracle's top-1 / top-2 margin is at most TWO bf16 ulps of the row's largest logit (the
    reference's logits are bf16 numbers and two faithful implementations sit one rounding step apart on each: which of
    such a pair argmax returns is decided by an accumulation order, not by the algorithm; the unmodified reference and
    the CPU oracle part at 0 / 1 / 2-ulp margins themselves) - and at least 80 % are reproduced exactly on the peaked-logit decoder
    (`synth.peaked_unembed_`, what a trained model looks like; measured 87.5 % = 14 of 16 over 128 steps, the two
    departures at margins of 0 and 2 ulps; the oracle's own bf16 mode reproduces 15 of its 16
    fp32 sequences).  On random-init weights the logits are nearly flat (mean margin 0.24 at max|logit| 2.5) and even
    the oracle's own bf16 mode reproduces only 25 % of its fp32 sequences, all at sub-ulp margins: there the exact rate
    is reported, not gated;
  * fp32 (configs[0]): one 256 x 1024 system through the full-size model, 64 greedy tokens, token-exact, latent 1e-3.
"""
import json
import os

import pytest
import torch

from acai_omr_b200 import decode_engine, synth
from acai_omr_b200.models import models as M
from oracle import parity as P
from oracle import vitomr_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
VOCAB = M.default_vocab_path()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENC_CFG = dict(patch_size=16, pe_max_height=60, pe_max_width=200, num_heads=12, interpolate=True, stacks=["fine_tune_blocks"])


def sub(sd, prefix):
    return {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}


def host_threads():
    return max(1, len(os.sched_getaffinity(0)))


def report(name, obj):
    """Parity numbers are kept next to the test result (gpurun_out/ travels back from the GPU box)."""
    print(name, json.dumps(obj))
    out = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_fullsize.jsonl"), "a") as f:
            f.write(json.dumps({"test": name, **obj}) + "\n")
    except OSError:
        pass


def full_decoder(batch, peaked, cache_dtype=torch.bfloat16, seed=2):
    torch.manual_seed(0)
    dec = M.OMRDecoder(1536, VOCAB, num_layers=12, hidden_dim=1024, num_heads=16, mlp_dim=4096, use_caching=True,
                       max_batch_size=batch, cache_dtype=cache_dtype)
    synth.rerandomise_(dec, seed)
    if peaked:
        synth.peaked_unembed_(dec)
    return dec.eval()


def memory_batch(batch, seed=5):
    lens = [(h // 16) * (w // 16) for h, w in synth.system_shapes(batch)]
    S = max(lens)
    g = torch.Generator().manual_seed(seed)
    mem = torch.randn(batch, S, 1024, generator=g)
    mask = torch.arange(S)[None, :] >= torch.tensor(lens)[:, None]
    return mem, mask, lens


def engine_for(vit, mem, mask, max_len):
    dec, dt = vit.decoder, torch.bfloat16
    packed, mp = dec._memory_packed(mem, mask, dt)
    eng = decode_engine.get_engine(vit, dec, mem.shape[0], max_len, mp.total, dt, mem.device)
    eng.prefill(packed, mp)
    eng.set_sampling(None)
    return eng


# ------------------------------------------------------------------------------------------------------------------
# 1. the engine against the UNMODIFIED reference's recorded bf16 run (head_dim 64, B = 6 -> TMA attention, split-K)
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["flat", "peaked"])
def test_bf16_engine_vs_reference_recorded_run(name):
    g = torch.load(os.path.join(ROOT, "tests", "golden", "golden_bf16.pt"), map_location="cpu", weights_only=False)[name]
    c = g["cfg"]
    enc = M.FineTuneOMREncoder(16, 4, 8, 1, num_layers=1, hidden_dim=c["enc_dim"], num_heads=2, mlp_dim=128)
    dec = M.OMRDecoder(c["max_lmx_seq_len"], VOCAB, num_layers=c["num_layers"], hidden_dim=c["hidden_dim"],
                       num_heads=c["num_heads"], mlp_dim=c["mlp_dim"], use_caching=True, max_batch_size=c["batch"],
                       cache_dtype=torch.bfloat16)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=c["head_dim"])
    missing, unexpected = vit.load_state_dict(g["sd"], strict=False)
    assert not unexpected and all(k.startswith("encoder.") for k in missing)
    vit = vit.to(DEV).eval()
    lat, mask = g["latent"].to(DEV), g["mask"].to(DEV)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        head = vit.transition_head(lat)
        assert decode_engine.eligible(vit.decoder, torch.bfloat16)
        eng = engine_for(vit, head, mask, c["max_len"])
        forced = eng.forced_logits(g["seqs"])
        seqs, logp, smask = vit.cached_greedy_generate(head, mask, max_len=c["max_len"])
    ref = g["logits"].float()[:, : forced.shape[1]]
    valid = g["seq_mask"][:, : forced.shape[1]]            # steps whose input prefix is part of the reference's sequence
    e_head = P.errors(head, g["head"], ~g["mask"])
    e = P.errors(forced, ref, valid)
    a = P.argmax_agreement(forced, ref, P.BF16_TOL, valid)
    d = P.divergence_report(seqs, g["seqs"], g["logits"].float(), P.BF16_TOL)
    report(f"engine_vs_reference_bf16[{name}]", {"head": e_head, "forced_logits": e, "argmax": a, "greedy": d})
    assert e_head["max_rel"] < P.BF16_TOL
    assert e["max_rel"] < P.BF16_TOL and e["l2_rel"] < P.BF16_TOL
    assert a["not_tie_disagreements"] == 0, a
    assert d["exact_or_within_bf16_resolution_rate"] == 1.0, d
    # log-probs of the tokens both sides chose, while the prefixes still agree: within the logit tolerance
    w = min(seqs.shape[1], g["seqs"].shape[1])
    prefix_same = (seqs.cpu()[:, :w] == g["seqs"][:, :w]).int().cumprod(dim=1).bool() & g["seq_mask"][:, :w]
    assert (logp.cpu()[:, :w] - g["log_probs"][:, :w].float())[prefix_same].abs().max() < P.BF16_TOL * e["ref_max"]


# ------------------------------------------------------------------------------------------------------------------
# 2. full-size decoder, bf16 engine, 128 steps against the live oracle
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("peaked", [False, True], ids=["random_init", "peaked"])
def test_fullsize_bf16_decode_engine_vs_oracle(peaked):
    B, max_len = (16 if peaked else 8), 129
    dec = full_decoder(B, peaked)
    sd = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    mem, mask, lens = memory_batch(B)
    mem = mem.bfloat16().float()                          # what the bf16 transition head hands over
    torch.set_num_threads(host_threads())
    with torch.no_grad():
        with O.autocast_bf16():
            seqs_b, logp_b, _, lg_b = O.cached_greedy_generate(mem, sd, 16, mask, max_len=max_len)
        seqs_f, _, _, lg_f = O.cached_greedy_generate(mem, sd, 16, mask, max_len=max_len)
        lg_b, lg_f = torch.stack(lg_b, 1).float(), torch.stack(lg_f, 1)
        forced_f = O.cached_forced_logits(mem, sd, 16, seqs_b, mask)        # fp32 oracle on the bf16 oracle's tokens
    vit = M.ViTOMR(torch.nn.Identity(), torch.nn.Identity(), dec).to(DEV).eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        eng = engine_for(vit, mem.to(DEV), mask.to(DEV), max_len)
        assert eng.use_splitk
        forced = eng.forced_logits(seqs_b)
        seqs, logp, smask = vit.cached_greedy_generate(mem.to(DEV), mask.to(DEV), max_len=max_len)
    n = min(forced.shape[1], lg_b.shape[1])
    forced = forced[:, :n]
    valid = O.create_inference_mask(seqs_b, dec.eos_idx)[:, :n]
    out = {"forced_vs_oracle_bf16": P.errors(forced, lg_b[:, :n], valid),
           "forced_vs_oracle_fp32": P.errors(forced, forced_f[:, :n], valid),
           "argmax_vs_oracle_bf16": P.argmax_agreement(forced, lg_b[:, :n], P.BF16_TOL, valid),
           "greedy_vs_oracle_bf16": P.divergence_report(seqs, seqs_b, lg_b, P.BF16_TOL),
           "greedy_vs_oracle_fp32": P.divergence_report(seqs, seqs_f, lg_f, P.BF16_TOL),
           "oracle_bf16_vs_oracle_fp32": P.divergence_report(seqs_b, seqs_f, lg_f, P.BF16_TOL),
           "steps": n, "batch": B, "memory_lens": lens, "launches_per_step": eng.launches_per_step}
    report(f"fullsize_decode_bf16[{'peaked' if peaked else 'random_init'}]", out)
    for key in ("forced_vs_oracle_bf16", "forced_vs_oracle_fp32"):
        assert out[key]["max_rel"] < P.BF16_TOL and out[key]["l2_rel"] < P.BF16_TOL, (key, out[key])
    assert out["argmax_vs_oracle_bf16"]["not_tie_disagreements"] == 0, out["argmax_vs_oracle_bf16"]
    for key in ("greedy_vs_oracle_bf16", "greedy_vs_oracle_fp32"):
        assert out[key]["exact_or_within_bf16_resolution_rate"] == 1.0, (key, out[key])
    if peaked:
        assert out["greedy_vs_oracle_bf16"]["exact_rate"] >= 0.8, out["greedy_vs_oracle_bf16"]
        assert out["greedy_vs_oracle_fp32"]["exact_rate"] >= 0.75, out["greedy_vs_oracle_fp32"]
        assert len({tuple(r.tolist()) for r in seqs_b}) > 1            # the memory decides the sequence


# ------------------------------------------------------------------------------------------------------------------
# 3. full-size encoder on config-2 systems, bf16 autocast and fp32, against the fp32 oracle
# ------------------------------------------------------------------------------------------------------------------
def test_fullsize_encoder_vs_oracle():
    torch.manual_seed(0)
    enc = M.FineTuneOMREncoder(16, 60, 200, 12, num_layers=12, hidden_dim=768, num_heads=12, mlp_dim=3072)
    synth.rerandomise_(enc, 1)
    enc = enc.eval()
    sd = {k: v.detach().clone() for k, v in enc.state_dict().items()}
    shapes = [synth.system_shapes(8)[i] for i in (0, 1, 3, 7)]              # 512x512, 352x704, 256x1024, 176x1408
    imgs = [synth.synth_system(h, w, i) for i, (h, w) in enumerate(shapes)]
    torch.set_num_threads(host_threads())
    with torch.no_grad():
        ref, mask_o = O.encoder_forward(imgs, sd, ENC_CFG)
    enc = enc.to(DEV)
    dev_imgs = [t.to(DEV) for t in imgs]
    with torch.no_grad():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out_b, mask = enc(dev_imgs)
        out_f, _ = enc(dev_imgs[2:3])                                     # fp32 mode: the 256 x 1024 system
    assert torch.equal(mask.cpu(), mask_o)
    valid = ~mask_o
    eb = P.errors(out_b, ref, valid)
    ef = P.errors(out_f[0], ref[2, : out_f.shape[1]])
    pad = P.errors(out_b.cpu()[mask_o], ref[mask_o]) if mask_o.any() else None      # pad rows = final-norm bias
    report("fullsize_encoder", {"bf16_vs_oracle": eb, "fp32_vs_oracle": ef, "pad_rows": pad,
                                "tokens": int(valid.sum()), "shapes": shapes})
    assert eb["max_rel"] < P.BF16_TOL and eb["l2_rel"] < P.BF16_TOL, eb
    assert ef["max_rel"] < P.FP32_TOL and ef["l2_rel"] < P.FP32_TOL, ef
    if pad is not None:
        assert pad["max_abs"] < 1e-6


# ------------------------------------------------------------------------------------------------------------------
# 4. configs[0]: one 256 x 1024 system, full-size model, fp32, greedy, token-exact
# ------------------------------------------------------------------------------------------------------------------
def test_config0_fp32_token_exact():
    torch.manual_seed(0)
    enc = M.FineTuneOMREncoder(16, 60, 200, 12, num_layers=12, hidden_dim=768, num_heads=12, mlp_dim=3072)
    dec = M.OMRDecoder(1536, VOCAB, num_layers=12, hidden_dim=1024, num_heads=16, mlp_dim=4096, use_caching=True,
                       max_batch_size=1, cache_dtype=torch.float32)
    vit = M.TeacherForcedViTOMR(enc, None, dec)
    synth.rerandomise_(vit, 1)
    synth.peaked_unembed_(vit.decoder)        # fp32 token-exactness needs margins above fp32 noise, not above bf16 noise
    vit = vit.eval()
    sd = {k: v.detach().clone() for k, v in vit.state_dict().items()}
    img = synth.synth_system(256, 1024, 0)
    max_len = 65
    torch.set_num_threads(host_threads())
    with torch.no_grad():
        lat_o, mask_o = O.encoder_forward([img], sub(sd, "encoder."), ENC_CFG)
        head_o = O.transition_head(lat_o, sd)
        seqs_o, logp_o, _, lg_o = O.cached_greedy_generate(head_o, sub(sd, "decoder."), 16, mask_o, max_len=max_len)
    vit = vit.to(DEV)
    with torch.no_grad():
        lat, mask = vit.encoder(img.to(DEV))                              # a bare (1, H, W) tensor, vitomr_inference.py:81
        head = vit.transition_head(lat)
        seqs, logp, _ = vit.cached_greedy_generate(head, mask, max_len=max_len)
    e_lat, e_head = P.errors(lat, lat_o), P.errors(head, head_o)
    d = P.divergence_report(seqs, seqs_o, torch.stack(lg_o, 1), P.FP32_TOL)
    report("config0_fp32", {"latent": e_lat, "head": e_head, "greedy": d, "tokens": int(seqs_o.shape[1])})
    assert e_lat["max_rel"] < P.FP32_TOL and e_head["max_rel"] < P.FP32_TOL
    assert torch.equal(seqs.cpu(), seqs_o), d
    assert torch.allclose(logp.cpu(), logp_o, atol=1e-3, rtol=1e-3)


# ------------------------------------------------------------------------------------------------------------------
# 5. the captured graph must notice weight updates
# ------------------------------------------------------------------------------------------------------------------
def test_engine_recaptures_after_weight_update():
    torch.manual_seed(0)
    mk = lambda: M.OMRDecoder(48, VOCAB, num_layers=2, hidden_dim=256, num_heads=4, mlp_dim=512, use_caching=True,  # noqa: E731
                              max_batch_size=8, cache_dtype=torch.bfloat16)
    dec = synth.peaked_unembed_(synth.rerandomise_(mk(), 3))
    other = synth.peaked_unembed_(synth.rerandomise_(mk(), 4), seed=9)
    vit = M.ViTOMR(torch.nn.Identity(), torch.nn.Identity(), dec).to(DEV).eval()
    g = torch.Generator().manual_seed(1)
    lat = torch.randn(8, 30, 256, generator=g).to(DEV)

    def run(v):
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            s, lp, _ = v.cached_greedy_generate(lat, None, max_len=32)
        return s.cpu(), lp.cpu()
    first = run(vit)
    assert torch.equal(run(vit)[0], first[0])                              # replay of the same graph
    vit.decoder.load_state_dict(other.state_dict())                        # version bump, new bf16 operand copies
    again = run(vit)
    fresh = M.ViTOMR(torch.nn.Identity(), torch.nn.Identity(), other).to(DEV).eval()
    want = run(fresh)
    assert torch.equal(again[0], want[0]) and torch.allclose(again[1], want[1], atol=1e-6)
    assert not torch.equal(again[0], first[0])
    with torch.no_grad():                                                  # an optimiser-style in-place update through torch
        vit.decoder.unembed.weight.mul_(-1.0)
        vit.decoder.decoder_blocks.layers[0].linear1.weight.mul_(1.5)
    third = run(vit)
    fresh2 = mk()
    fresh2.load_state_dict(vit.decoder.state_dict())
    want2 = run(M.ViTOMR(torch.nn.Identity(), torch.nn.Identity(), fresh2).to(DEV).eval())
    assert torch.equal(third[0], want2[0]) and not torch.equal(third[0], again[0])
