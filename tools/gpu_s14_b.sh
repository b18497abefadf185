# session 14, call B: chunked host frames (tests + e2e sweep over chunk counts)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_s14b.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_s14b.log
for c in 1 2 3 4 6 8; do
  timeout 300 python bench.py --steps 30 --warmup 3 --host-chunks $c --no-cpu-baseline 2>gpurun_out/bench_s14b_c$c.err | tail -1 > gpurun_out/bench_s14b_c$c.json
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_s14b_c$c.json")); print("chunks $c", "value ms", round(d["ms_per_step"],3), "e2e ms", round(d["e2e"]["ms_per_step"],3), d["e2e"]["path"][:60])
except Exception as e: print("no json", e); print(open("gpurun_out/bench_s14b_c$c.err").read()[-1500:])
PY
done
