timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>gpurun_out/err.log | tail -1 | tee gpurun_out/bench_r1_g.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['breakdown_ms'], d['gpu_launches'])" || tail gpurun_out/err.log
