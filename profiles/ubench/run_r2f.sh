mkdir -p gpurun_out
F="ms/step\|gelu_bwd\|,res\]\|linear_dropout"
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== tf coalesced residual loads"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "$F" | head -14
echo "== tf ACAI_GEMM_RES_COALESCED=0"; ACAI_GEMM_RES_COALESCED=0 timeout 600 python profiles/prof_train.py tf 2>&1 | grep "$F" | head -14
echo "== mae"; timeout 600 python profiles/prof_train.py mae 2>&1 | grep "ms/step" | head -2
echo "== mae ACAI_GEMM_RES_COALESCED=0"; ACAI_GEMM_RES_COALESCED=0 timeout 600 python profiles/prof_train.py mae 2>&1 | grep "ms/step" | head -2
} > gpurun_out/r2h_c.log 2>&1
tail -50 gpurun_out/r2h_c.log
