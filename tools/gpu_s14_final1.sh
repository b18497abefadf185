# session 14, final N=1 lines (with cpu_baseline), both headline scenes + large world, reference arm (short)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 50 --warmup 5 2>gpurun_out/bench_final_n1.err | tail -1 > gpurun_out/bench_final_n1.json; cut -c1-300 gpurun_out/bench_final_n1.json; echo
timeout 600 python bench.py --steps 50 --warmup 5 --scene world8_stress --no-cpu-baseline 2>gpurun_out/bench_final_stress_n1.err | tail -1 > gpurun_out/bench_final_stress_n1.json
timeout 600 python bench.py --steps 20 --warmup 3 --grid-size 256 --no-cpu-baseline 2>gpurun_out/bench_final_g256_n1.err | tail -1 > gpurun_out/bench_final_g256_n1.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 2>gpurun_out/bench_final_ref.err | tail -1 > gpurun_out/bench_final_ref.json; cut -c1-400 gpurun_out/bench_final_ref.json; echo
python - <<PY
import json
for f in ("bench_final_n1","bench_final_stress_n1","bench_final_g256_n1"):
    try:
        d=json.load(open("gpurun_out/%s.json"%f)); print(f, "ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "e2e ms", round(d["e2e"]["ms_per_step"],3), round(d["e2e"]["value"]), d["breakdown_ms"], d["roofline"]["frac"], d["roofline"]["traffic"], d["roofline"]["l2"]["frac"], d.get("clocks"))
    except Exception as e: print(f, "no json", e); print(open("gpurun_out/%s.err"%f).read()[-1500:])
PY
