N=$1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 2>gpurun_out/err_n$N.log | tail -1 | tee gpurun_out/bench_r1_n$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], '| Mrays/s', round(d['value'],1), '| ms', round(d['ms_per_step'],3), '| e2e ms', round(d['e2e']['ms_per_step'],3), d['breakdown_ms'], d['clocks'])" || tail -5 gpurun_out/err_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --scene world8 2>gpurun_out/err_n$N.log | tail -1 | tee gpurun_out/bench_r1_world8_n$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('world8 N', d['n_gpus'], '| Mrays/s', round(d['value'],1), '| ms', round(d['ms_per_step'],3), '| e2e ms', round(d['e2e']['ms_per_step'],3), d['breakdown_ms'])" || tail -5 gpurun_out/err_n$N.log
