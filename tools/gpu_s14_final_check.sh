# session 14, final check at HEAD: whole GPU suite, smoke, N=1 bench line with cpu_baseline
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_final.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 50 --warmup 5 2>gpurun_out/bench_final2_n1.err | tail -1 > gpurun_out/bench_final2_n1.json
python - <<PY
import json
d=json.load(open("gpurun_out/bench_final2_n1.json")); print("ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "e2e ms", round(d["e2e"]["ms_per_step"],3), round(d["e2e"]["value"]), d["breakdown_ms"], "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"],4), d["roofline"]["traffic"], "l2", round(d["roofline"]["l2"]["frac"],3), d["cpu_baseline"]["value"], d["clocks"]["reasons"], d["gpu_launches"])
PY
