# last command of round 2h: the GPU test-suite and smoke() on the final build
mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
} > gpurun_out/r2h_g.log 2>&1
tail -12 gpurun_out/r2h_g.log
