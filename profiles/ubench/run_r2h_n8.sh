# round 2h, 8 GPUs: the bench's encoder + training + strong-scaling legs under torchrun (no decode / eager / config0 / CPU baseline)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29522 bench.py --gpus 8 --steps 10 --warmup 3 --no-decode --no-torch-eager --no-config0 --no-cpu-baseline > gpurun_out/r2h_bench_line_n8.json 2> gpurun_out/r2h_bench_n8.err; echo "bench n8 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2h_bench_line_n8.json").read().strip().splitlines()[-1])
t, m = d["train_teacher_forced"], d["train_mae"]
print("n_gpus", d["n_gpus"], "enc", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "tf", round(t["value"], 1), round(t["ms_per_step"], 2), t.get("collective"), "mae", round(m["value"], 1), round(m["ms_per_step"], 2), m.get("collective"), "strong", d.get("strong_scaling", {}).get("value"))
PY
