mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== bench"; timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/r2h_bench_line_n1.json 2> gpurun_out/r2h_bench.err; echo rc=$?
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2h_bench_line_n1.json").read().strip().splitlines()[-1])
s = d["config"]["secondary"]
print("value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "roofline", round(d["roofline"]["frac"], 3), "clocks", d["clocks"])
print({k: s[k] for k in ("decode_tokens_per_s", "decode_step_hbm_frac", "tf_ms_per_step", "tf_systems_per_s", "mae_ms_per_step", "mae_systems_per_s", "tf_gemm_frac")})
PY
} > gpurun_out/r2h_final.log 2>&1
tail -20 gpurun_out/r2h_final.log
