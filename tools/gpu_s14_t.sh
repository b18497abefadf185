# session 14, call T: RT_RENDER_SORT_RAYS — tests and A/B on the bounce-heavy scenes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_spheres.py tests/test_motion.py -m gpu -x -q > gpurun_out/pytest_gpu_s14t.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_s14t.log
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); b=d["breakdown_ms"]; print("$2", "ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "| shade %.3f waves %.3f (%d)" % (b["shade"], b["bounce_waves"], b["waves"]))
except Exception as e: print("no json", e); print(open("$1".replace(".json",".err")).read()[-1500:])
PY
}
for cfg in "world8_stress 9" "spheres8 5" "grid64_mixed 9"; do set -- $cfg
  for so in "" "--sort-rays"; do
    timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --scene $1 --depth $2 $so 2>gpurun_out/bench_s14t_$1$so.err | tail -1 > gpurun_out/bench_s14t_$1$so.json; show gpurun_out/bench_s14t_$1$so.json "$1 $so"
  done
done
