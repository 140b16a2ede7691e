mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
echo "== tf dgrad kn"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "ms/step\|acai_transpose\|acai_linear_dgrad\|sum " | head -24
echo "== tf ACAI_DGRAD_KN=0"; ACAI_DGRAD_KN=0 timeout 600 python profiles/prof_train.py tf 2>&1 | grep "ms/step\|acai_transpose\|sum " | head
echo "== mae dgrad kn"; timeout 600 python profiles/prof_train.py mae 2>&1 | grep "ms/step\|acai_transpose\|sum " | head
echo "== mae ACAI_DGRAD_KN=0"; ACAI_DGRAD_KN=0 timeout 600 python profiles/prof_train.py mae 2>&1 | grep "ms/step\|acai_transpose\|sum " | head
B="bench.py --steps 10 --warmup 3 --no-decode --no-torch-eager --no-strong --no-config0 --no-cpu-baseline"
echo "== bench train legs"; timeout 900 python $B > gpurun_out/r2f_bt1.json 2>/dev/null; echo rc=$?
echo "== bench train legs ACAI_GEMM_WAVE_QUANT=0"; ACAI_GEMM_WAVE_QUANT=0 timeout 900 python $B > gpurun_out/r2f_bt0.json 2>/dev/null; echo rc=$?
python - <<'PY'
import json
for f in ("r2f_bt1", "r2f_bt0"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    m, t = d["train_mae"], d["train_teacher_forced"]
    print(f, "enc", round(d["ms_per_step"], 3), "tf", round(t["ms_per_step"], 2), "mae", round(m["ms_per_step"], 3), {k: v for k, v in m["kernels_ms_per_step"].items() if k.startswith("acai_linear")})
PY
} > gpurun_out/r2f_h.log 2>&1
tail -60 gpurun_out/r2f_h.log
