for cfg in "world8_stress 3840 2160 2" "world8_stress 640 480 2" "world4 1920 1080 9" "world8 3840 2160 9" "world1 640 480 2" "grid64 3840 2160 2" "grid128 3840 2160 2" "grid64_mixed 1920 1080 4"; do
  set -- $cfg
  timeout 120 python bench.py --scene $1 --width $2 --height $3 --depth $4 --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['config']['workload'], '| Mrays/s', round(d['value'],1), '| ms', round(d['ms_per_step'],3), '| e2e ms', round(d['e2e']['ms_per_step'],3), '| rays', d['config']['rays_per_frame'], '| tris', d['config']['triangles'], d['breakdown_ms'])"
done 2>&1 | tee gpurun_out/bench_r1_configs.log
