# session 14, final scaling lines N=8,4,2 on one 8-GPU box (N=1 measured on a 1-GPU box: profiles/r1_bench_final_n1.json)
mkdir -p gpurun_out
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2954$n bench.py --gpus $n --steps 40 --warmup 5 2>gpurun_out/bench_final_n$n.err | tail -1 > gpurun_out/bench_final_n$n.json
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_final_n$n.json")); print("N=$n", "value ms", round(d["ms_per_step"],4), "Mrays/s", round(d["value"]), "| e2e ms", round(d["e2e"]["ms_per_step"],3), round(d["e2e"]["value"]), d["breakdown_ms"]["build"], d["breakdown_ms"]["trace"], d["clocks"])
except Exception as e: print("no json", e); print(open("gpurun_out/bench_final_n$n.err").read()[-1500:])
PY
done
