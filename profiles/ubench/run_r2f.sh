mkdir -p gpurun_out
F="ms/step\|gelu_train\|gelu_bwd\]\|gelu_bwd,\|linear_dropout\|acai_dropout\|,gelu_bwd"
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== tf fold2 hidden-dropout hash"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "$F" | head -12
echo "== tf mix32 (alt library)"; ACAI_LIB_PATH=$PWD/acai_omr_b200/csrc/build/alt/libacai_mix32.so timeout 600 python profiles/prof_train.py tf 2>&1 | grep "$F" | head -12
echo "== tf fold2 again"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "$F" | head -12
} > gpurun_out/r2f_i.log 2>&1
tail -50 gpurun_out/r2f_i.log
