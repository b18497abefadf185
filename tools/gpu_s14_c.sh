# session 14, call C (2 GPUs): geometric chunks at N=1, host bands at N=2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "chunked or bands or shared" 2>&1 | tail -3
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); print("$2", "n", d["n_gpus"], "value ms", round(d["ms_per_step"],3), "Mrays/s", round(d["value"]), "| e2e ms", round(d["e2e"]["ms_per_step"],3), "Mrays/s", round(d["e2e"]["value"]), d["e2e"]["path"][-40:])
except Exception as e: print("no json", e); print(open("$1".replace(".json",".err")).read()[-1500:])
PY
}
for c in 0 2 3; do
  timeout 300 python bench.py --steps 30 --warmup 3 --host-chunks $c --no-cpu-baseline 2>gpurun_out/bench_s14c_c$c.err | tail -1 > gpurun_out/bench_s14c_c$c.json; show gpurun_out/bench_s14c_c$c.json "chunks=$c"
done
for bh in 4 8 16; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 30 --warmup 3 --band-h $bh 2>gpurun_out/bench_s14c_n2_b$bh.err | tail -1 > gpurun_out/bench_s14c_n2_b$bh.json; show gpurun_out/bench_s14c_n2_b$bh.json "N=2 band_h=$bh"
done
