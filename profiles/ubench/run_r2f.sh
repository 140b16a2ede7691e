# round 2f: bit-parallel attention dropout mask - full GPU test-suite, isolated attention kernels, the teacher-forced step
mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -6
echo "== attention kernels"; timeout 600 python profiles/ubench/kernel_bench.py attn attn_bwd drop 2>&1 | tail -14
echo "== teacher-forced step"; timeout 600 python profiles/prof_train.py tf 2>&1 | tail -40
} > gpurun_out/r2f_b.log 2>&1
tail -70 gpurun_out/r2f_b.log
