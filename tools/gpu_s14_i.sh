# session 14, call I: textures + spheres on the GPU, whole GPU suite, perf check of the shading kernels after the kd plumbing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_s14i.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu_s14i.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_s14i.err | tail -1 > gpurun_out/bench_s14i.json
timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --scene world8_stress 2>gpurun_out/bench_s14i_stress.err | tail -1 > gpurun_out/bench_s14i_stress.json
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --scene textured8 --depth 2 2>gpurun_out/bench_s14i_tex.err | tail -1 > gpurun_out/bench_s14i_tex.json
python - <<PY
import json
for f in ("bench_s14i","bench_s14i_stress","bench_s14i_tex"):
    try:
        d=json.load(open("gpurun_out/%s.json"%f)); print(f, "ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "e2e ms", round(d["e2e"]["ms_per_step"],3), d["breakdown_ms"])
    except Exception as e: print(f, "no json", e); print(open("gpurun_out/%s.err"%f).read()[-1500:])
PY
