# session 14, call A: HEAD check + big-world first contact
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_s14a.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_s14a.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 50 --warmup 5 2>gpurun_out/bench_s14a_n1.err | tail -1 > gpurun_out/bench_s14a_n1.json; cut -c1-400 gpurun_out/bench_s14a_n1.json; echo
for sc in "grid64 0" "grid128 0" "world16 256"; do set -- $sc
  timeout 300 python bench.py --steps 10 --warmup 3 --scene $1 --grid-size $2 --no-cpu-baseline 2>gpurun_out/bench_s14a_$1_$2.err | tail -1 > gpurun_out/bench_s14a_$1_$2.json; echo "$1 $2 rc=$?"; cut -c1-300 gpurun_out/bench_s14a_$1_$2.json; echo; tail -2 gpurun_out/bench_s14a_$1_$2.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_s14a_$1_$2.json")); print(d["config"]["instances"], d["config"]["triangles"], d["breakdown_ms"], d["roofline"]["l2"], d["roofline"]["frac"])
except Exception as e: print("no json", e)
PY
done
