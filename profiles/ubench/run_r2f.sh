# the command most gpurun calls of rounds 2f - 2i were variations of: the GPU test-suite, smoke() and the A/B legs of one change
# (each change note under profiles/ names its own log).  Left here in its round-end form.
mkdir -p gpurun_out
{
echo "== gpu tests"; timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -3
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== teacher-forced / MAE steps"; timeout 600 python profiles/prof_train.py tf 2>&1 | grep "ms/step"; timeout 600 python profiles/prof_train.py mae 2>&1 | grep "ms/step"
} > gpurun_out/round_end.log 2>&1
tail -12 gpurun_out/round_end.log
