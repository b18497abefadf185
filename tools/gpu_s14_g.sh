# session 14, call G: per-digit radix scan — GPU tests (BVH bit-exact at 256 k and 1.3 M triangles) and big-world benches
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_s14g.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_s14g.log
for sc in "grid64 0" "grid128 0" "world16 256"; do set -- $sc
  timeout 300 python bench.py --steps 10 --warmup 3 --scene $1 --grid-size $2 --no-cpu-baseline 2>gpurun_out/bench_s14g_$1_$2.err | tail -1 > gpurun_out/bench_s14g_$1_$2.json
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_s14g_$1_$2.json")); print("$1 $2", d["config"]["triangles"], "ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "e2e ms", round(d["e2e"]["ms_per_step"],3), d["breakdown_ms"])
except Exception as e: print("no json", e); print(open("gpurun_out/bench_s14g_$1_$2.err").read()[-1500:])
PY
done
