mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== tf"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "ms/step\|layernorm" | head -5
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"layernorm_bwd_vec" --launch-skip 70 --launch-count 2 -f -o gpurun_out/r2h_lnbwd python profiles/prof_train.py tf > gpurun_out/r2h_ncu_ln.log 2>&1; echo "ncu rc=$?"
} > gpurun_out/r2h_b.log 2>&1
tail -12 gpurun_out/r2h_b.log
