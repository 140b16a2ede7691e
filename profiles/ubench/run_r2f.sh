mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
ACAI_WGRAD_SPLIT=1 timeout 300 python profiles/ubench/wgrad_precision.py 2>&1 | tail -8
echo "== tf"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "ms/step\|acai_wgrad\|x227\|x256\]" | head -20
echo "== bench mae/tf legs"; timeout 900 python bench.py --steps 10 --warmup 3 --no-decode --no-torch-eager --no-strong --no-config0 --no-cpu-baseline > gpurun_out/r2f_bench_train.json 2> gpurun_out/r2f_bench_train.err; echo rc=$?
} > gpurun_out/r2f_g.log 2>&1
tail -60 gpurun_out/r2f_g.log
