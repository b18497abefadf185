# session 14, call M (8 GPUs): own mailbox barrier vs NCCL all-reduce barrier at N=2,8; signalled push test
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "signalled or directly" 2>&1 | tail -3
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); print("$2", "n", d["n_gpus"], "value ms", round(d["ms_per_step"],4), "Mrays/s", round(d["value"]), "| e2e ms", round(d["e2e"]["ms_per_step"],3), d["breakdown_ms"]["build"], d["breakdown_ms"]["trace"])
except Exception as e: print("no json", e); print(open("$1".replace(".json",".err")).read()[-1500:])
PY
}
for n in 8 2; do for g in p2p p2p-nccl p2p; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 40 --warmup 5 --gather $g 2>gpurun_out/bench_s14m_n${n}_$g.err | tail -1 > gpurun_out/bench_s14m_n${n}_$g.json; show gpurun_out/bench_s14m_n${n}_$g.json "N=$n $g"
done; done
