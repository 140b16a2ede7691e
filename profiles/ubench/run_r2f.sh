# round 2f: bias gradients inside the weight-gradient kernel - full GPU test-suite, teacher-forced and MAE steps with / without
mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -6
echo "== teacher-forced step, fused bias gradients"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep -v "^  acai_linear\|^  acai_gelu\|^  acai_un\|^  acai_pa\|^  acai_em\|^  acai_sc\|^  acai_ce" | head -60
echo "== teacher-forced step, ACAI_FUSE_BIAS_GRAD=0"; ACAI_FUSE_BIAS_GRAD=0 timeout 600 python profiles/prof_train.py tf 2>&1 | grep -v "^  acai_linear\|^  acai_gelu\|^  acai_un\|^  acai_pa\|^  acai_em\|^  acai_sc\|^  acai_ce" | head -60
echo "== MAE step, fused"; timeout 600 python profiles/prof_train.py mae 2>&1 | head -30
echo "== MAE step, ACAI_FUSE_BIAS_GRAD=0"; ACAI_FUSE_BIAS_GRAD=0 timeout 600 python profiles/prof_train.py mae 2>&1 | head -30
} > gpurun_out/r2f_c.log 2>&1
tail -150 gpurun_out/r2f_c.log
