# session 14, call D (8 GPUs): host bands at N=8 and N=4
mkdir -p gpurun_out
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); print("$2", "n", d["n_gpus"], "value ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "| e2e ms", round(d["e2e"]["ms_per_step"],3), "Mrays/s", round(d["e2e"]["value"]), d["breakdown_ms"]["build"], d["breakdown_ms"]["trace"])
except Exception as e: print("no json", e); print(open("$1".replace(".json",".err")).read()[-1500:])
PY
}
for n in 8 4; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 30 --warmup 3 2>gpurun_out/bench_s14d_n$n.err | tail -1 > gpurun_out/bench_s14d_n$n.json; show gpurun_out/bench_s14d_n$n.json "N=$n"
done
