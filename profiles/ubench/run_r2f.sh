mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== tf TMA-ring LayerNorm backward"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "ms/step\|layernorm" | head -4
echo "== tf ACAI_LN_BWD_TMA=0"; ACAI_LN_BWD_TMA=0 timeout 600 python profiles/prof_train.py tf 2>&1 | grep "ms/step\|layernorm" | head -4
echo "== mae"; timeout 600 python profiles/prof_train.py mae 2>&1 | grep "ms/step\|layernorm" | head -4
echo "== mae ACAI_LN_BWD_TMA=0"; ACAI_LN_BWD_TMA=0 timeout 600 python profiles/prof_train.py mae 2>&1 | grep "ms/step\|layernorm" | head -4
} > gpurun_out/r2h_d.log 2>&1
tail -30 gpurun_out/r2h_d.log
