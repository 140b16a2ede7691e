"""The bf16 greedy-token gate on a larger sample than tests/test_fullsize_gpu.py uses: full-size peaked decoder, B sequences,
128 steps, engine vs the oracle's bf16 mode and vs the fp32 oracle, as a function of how peaked the next-token distributions
are (synth.peaked_unembed_'s c: 0.05 is what the tests use).  python profiles/ubench/token_gate_sample.py [B] [c ...]"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_fullsize_gpu as T  # noqa: E402
from oracle import parity as P  # noqa: E402
from oracle import vitomr_oracle as O  # noqa: E402
from acai_omr_b200.models import models as M  # noqa: E402

from acai_omr_b200 import synth  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
CS = [float(a) for a in sys.argv[2:]] or [0.05]
max_len = 129
results = []
for c_peak in CS:
  dec = T.full_decoder(B, False)
  synth.peaked_unembed_(dec, c=c_peak)
  sd = {k: v.detach().clone() for k, v in dec.state_dict().items()}
  mem, mask, lens = T.memory_batch(B)
  mem = mem.bfloat16().float()
  torch.set_num_threads(T.host_threads())
  t0 = time.time()
  with torch.no_grad():
      with O.autocast_bf16():
          seqs_b, _, _, lg_b = O.cached_greedy_generate(mem, sd, 16, mask, max_len=max_len)
      seqs_f, _, _, lg_f = O.cached_greedy_generate(mem, sd, 16, mask, max_len=max_len)
  lg_b, lg_f = torch.stack(lg_b, 1).float(), torch.stack(lg_f, 1)
  t_oracle = time.time() - t0
  vit = M.ViTOMR(torch.nn.Identity(), torch.nn.Identity(), dec).to("cuda").eval()
  with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
      seqs, logp, smask = vit.cached_greedy_generate(mem.to("cuda"), mask.to("cuda"), max_len=max_len)
  out = {"c": c_peak, "batch": B, "steps": max_len - 1, "oracle_seconds": round(t_oracle, 1),
         "engine_vs_oracle_bf16": P.divergence_report(seqs, seqs_b, lg_b, P.BF16_TOL),
         "engine_vs_oracle_fp32": P.divergence_report(seqs, seqs_f, lg_f, P.BF16_TOL),
         "oracle_bf16_vs_oracle_fp32": P.divergence_report(seqs_b, seqs_f, lg_f, P.BF16_TOL)}
  top2 = lg_f.topk(2, dim=-1).values
  out["fp32_oracle_mean_margin"] = float((top2[..., 0] - top2[..., 1]).mean())
  out["fp32_oracle_max_logit"] = float(lg_f.abs().max())
  print("c =", c_peak, "mean top-1/top-2 margin", round(out["fp32_oracle_mean_margin"], 2), "max |logit|", round(out["fp32_oracle_max_logit"], 1))
  results.append(out)
  for k in ("engine_vs_oracle_bf16", "engine_vs_oracle_fp32", "oracle_bf16_vs_oracle_fp32"):
      d = out[k]
      print(k, "exact", d["exact"], "/", d["sequences"], "exact_or_within_bf16_resolution", d["exact_or_within_bf16_resolution_rate"],
            "margins(ulps)", [round(r["ref_margin"] / r["bf16_ulp"], 2) for r in d["divergences"]])
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(results, open(os.path.join(ROOT, "gpurun_out", f"token_gate_B{B}.json"), "w"))
