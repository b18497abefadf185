# session 14, call U: k_shadow at 4 blocks per SM (64 registers, small spills) against 3 (80 registers)
mkdir -p gpurun_out
cp gpu-ray-tracer_b200/librtb200.so /tmp/librtb200_base.so
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); b=d["breakdown_ms"]; print("$2", "ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "| shadow %.3f shade %.3f" % (b["k_shadow"], b["shade"]))
except Exception as e: print("no json", e); print(open("$1".replace(".json",".err")).read()[-800:])
PY
}
for v in base mb4 base mb4; do
  if [ $v = base ]; then cp /tmp/librtb200_base.so gpu-ray-tracer_b200/librtb200.so; else cp tools/variants/librtb200_$v.so gpu-ray-tracer_b200/librtb200.so; fi
  timeout 120 python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_s14u_$v.err | tail -1 > gpurun_out/bench_s14u_$v.json; show gpurun_out/bench_s14u_$v.json "world16 $v"
done
cp /tmp/librtb200_base.so gpu-ray-tracer_b200/librtb200.so
