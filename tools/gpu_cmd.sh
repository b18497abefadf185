timeout 300 python -m pytest tests -m gpu -q 2>&1 | tail -3
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | tee gpurun_out/bench_r1_f.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['breakdown_ms'])"
python tools/profile_frame.py world16 3840 2160 9 3 | tail -1
python tools/profile_frame.py world16 3840 2160 9 3 65 | tail -1
