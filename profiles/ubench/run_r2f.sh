mkdir -p gpurun_out
{
for sg in 0 3000 6000 9000 12000 0; do
echo "== ACAI_GLN_STAGGER=$sg"; ACAI_GLN_STAGGER=$sg LNB_SHAPE=63664,768,768 timeout 300 python profiles/ubench/linear_ln_bench.py 2>&1 | grep "res=float32" | cut -c1-110
ACAI_GLN_STAGGER=$sg LNB_SHAPE=63664,768,3072 timeout 300 python profiles/ubench/linear_ln_bench.py 2>&1 | grep "res=float32" | cut -c1-110
done
} > gpurun_out/r2i_a.log 2>&1
tail -30 gpurun_out/r2i_a.log
