# last command of round 2: the driver's bench command on the final build + the ncu launch list of its encoder leg
mkdir -p gpurun_out
timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/r2i_bench_line_n1.json 2> gpurun_out/r2i_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2i_bench_line_n1.json").read().strip().splitlines()[-1])
s = d["config"]["secondary"]
print("value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "roofline", round(d["roofline"]["frac"], 3), "clocks", d["clocks"])
print(d["kernels_ms_per_step"])
print({k: s[k] for k in ("decode_tokens_per_s", "decode_step_hbm_frac", "tf_ms_per_step", "tf_systems_per_s", "mae_ms_per_step", "mae_systems_per_s", "tf_gemm_frac")})
PY
ENC="bench.py --steps 2 --warmup 1 --no-decode --no-train --no-torch-eager --no-strong --no-config0 --no-cpu-baseline"
timeout 600 python $ENC > gpurun_out/r2i_enc_short.json 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2i_encoder_launches.csv python $ENC > gpurun_out/r2i_ncu_enc.log 2>&1; echo "enc ncu rc=$?"
