mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== decode, page table read once per chunk"; timeout 600 python profiles/prof_decode.py 256 32 1 2>&1 | tail -3
echo "== decode, old library"; ACAI_LIB_PATH=$PWD/acai_omr_b200/csrc/build/alt/libacai_tbl_old.so timeout 600 python profiles/prof_decode.py 256 32 1 2>&1 | tail -3
echo "== decode again (new)"; timeout 600 python profiles/prof_decode.py 256 2>&1 | tail -1
} > gpurun_out/r2h_a.log 2>&1
tail -30 gpurun_out/r2h_a.log
