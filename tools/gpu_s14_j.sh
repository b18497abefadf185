# session 14, call J (2 GPUs): kernels write the pinned host canvas directly — tests, A/B at N=1, N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_s14j.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_s14j.log
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); print("$2", "n", d["n_gpus"], "value ms", round(d["ms_per_step"],3), "| e2e ms", round(d["e2e"]["ms_per_step"],3), "Mrays/s", round(d["e2e"]["value"]), d["e2e"]["path"][:48])
except Exception as e: print("no json", e); print(open("$1".replace(".json",".err")).read()[-1500:])
PY
}
for v in "" "--host-staged"; do
  timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline $v 2>gpurun_out/bench_s14j_n1$v.err | tail -1 > gpurun_out/bench_s14j_n1$v.json; show gpurun_out/bench_s14j_n1$v.json "N=1 $v"
  timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --scene world8_stress $v 2>gpurun_out/bench_s14j_st$v.err | tail -1 > gpurun_out/bench_s14j_st$v.json; show gpurun_out/bench_s14j_st$v.json "stress N=1 $v"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 30 --warmup 3 $v 2>gpurun_out/bench_s14j_n2$v.err | tail -1 > gpurun_out/bench_s14j_n2$v.json; show gpurun_out/bench_s14j_n2$v.json "N=2 $v"
done
