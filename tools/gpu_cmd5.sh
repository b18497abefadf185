timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
for cfg in "world16 3840 2160 9" "world8_stress 3840 2160 2" "world1 640 480 2"; do
  set -- $cfg
  timeout 120 python bench.py --scene $1 --width $2 --height $3 --depth $4 --steps 20 --warmup 3 --no-cpu-baseline 2>gpurun_out/err.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['config']['workload'], '| Mrays/s', round(d['value'],1), '| ms', round(d['ms_per_step'],3), '| e2e ms', round(d['e2e']['ms_per_step'],3), '| rays', d['config']['rays_per_frame'], 'culled', round(d['config']['culled_frac'],3), d['breakdown_ms'])" || tail -5 gpurun_out/err.log
done 2>&1 | tee gpurun_out/bench_r1_configs3.log
