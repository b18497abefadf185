# session 14, call P: dynamic work distribution in k_shadow — parity tests + A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_spheres.py tests/test_texture.py tests/test_motion.py -m gpu -x -q > gpurun_out/pytest_gpu_s14p.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_s14p.log
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); b=d["breakdown_ms"]; print("$2", "ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "| e2e", round(d["e2e"]["ms_per_step"],3), "| primary %.3f shade %.3f shadow %.3f waves %.3f" % (b["k_primary"], b["shade"], b["k_shadow"], b["bounce_waves"]))
except Exception as e: print("no json", e); print(open("$1".replace(".json",".err")).read()[-1500:])
PY
}
for cfg in "world16 0" "world8_stress 0" "world16 256"; do set -- $cfg
  for st in 0 1; do
    RT_SHADOW_STATIC=$st timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --scene $1 --grid-size $2 2>gpurun_out/bench_s14p_$1_$2_s$st.err | tail -1 > gpurun_out/bench_s14p_$1_$2_s$st.json; show gpurun_out/bench_s14p_$1_$2_s$st.json "$1 grid $2 static=$st"
  done
done
