# round 2f, 2 GPUs: data-parallel equality check + the bench's training legs under torchrun
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 profiles/ubench/dp_overlap_check.py > gpurun_out/r2f_n2_dp_check.log 2>&1; echo "dp check rc=$?"; tail -6 gpurun_out/r2f_n2_dp_check.log
timeout 900 $TR --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-decode --no-torch-eager --no-config0 --no-cpu-baseline > gpurun_out/r2f_bench_line_n2.json 2> gpurun_out/r2f_bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2f_bench_line_n2.json").read().strip().splitlines()[-1])
t, m = d["train_teacher_forced"], d["train_mae"]
print("n_gpus", d["n_gpus"], "enc", d["value"], "tf", t["value"], t["ms_per_step"], t.get("collective"), "mae", m["value"], m["ms_per_step"], "strong", d.get("strong_scaling", {}).get("value"))
PY
