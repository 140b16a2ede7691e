mkdir -p gpurun_out
F="ms/step\|gelu_bwd"
{
echo "== ffn tests with ACAI_GEMM_EPI16=3"; ACAI_GEMM_EPI16=3 timeout 900 python -m pytest tests -q -m gpu -x -k "gelu or feed_forward or dgrad or teacher_forced or mae" 2>&1 | tail -3
echo "== tf default (1)"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "$F" | head -8
echo "== tf ACAI_GEMM_EPI16=3"; ACAI_GEMM_EPI16=3 timeout 600 python profiles/prof_train.py tf 2>&1 | grep "$F" | head -8
echo "== mae default"; timeout 600 python profiles/prof_train.py mae 2>&1 | grep "$F" | head -8
echo "== mae ACAI_GEMM_EPI16=3"; ACAI_GEMM_EPI16=3 timeout 600 python profiles/prof_train.py mae 2>&1 | grep "$F" | head -8
} > gpurun_out/r2g_c.log 2>&1
tail -50 gpurun_out/r2g_c.log
