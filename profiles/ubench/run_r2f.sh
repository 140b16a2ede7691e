# round 2f: residual-gradient hand-over (GradBox), NVTX ranges - full GPU test-suite, teacher-forced step with / without
mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -6
echo "== teacher-forced step, GradBox"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep -v "^  acai_linear\|^  acai_gelu\|^  acai_un\|^  acai_pa\|^  acai_em\|^  acai_sc\|^  acai_ce\|^  acai_wgrad" | head -30
echo "== teacher-forced step, ACAI_GRAD_BOX=0"; ACAI_GRAD_BOX=0 timeout 600 python profiles/prof_train.py tf 2>&1 | grep -v "^  acai_linear\|^  acai_gelu\|^  acai_un\|^  acai_pa\|^  acai_em\|^  acai_sc\|^  acai_ce\|^  acai_wgrad" | head -30
} > gpurun_out/r2f_d.log 2>&1
tail -80 gpurun_out/r2f_d.log
