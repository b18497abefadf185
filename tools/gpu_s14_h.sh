# session 14, call H: spheres on the GPU + whole GPU suite + bench sanity (perf must not move) + e2e on the big world
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_s14h.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu_s14h.log
timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_s14h.err | tail -1 > gpurun_out/bench_s14h.json
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --grid-size 256 2>gpurun_out/bench_s14h_g256.err | tail -1 > gpurun_out/bench_s14h_g256.json
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --scene spheres8 --depth 5 2>gpurun_out/bench_s14h_sph.err | tail -1 > gpurun_out/bench_s14h_sph.json
python - <<PY
import json
for f in ("bench_s14h","bench_s14h_g256","bench_s14h_sph"):
    try:
        d=json.load(open("gpurun_out/%s.json"%f)); print(f, "ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "e2e ms", round(d["e2e"]["ms_per_step"],3), d["breakdown_ms"])
    except Exception as e: print(f, "no json", e); print(open("gpurun_out/%s.err"%f).read()[-1500:])
PY
