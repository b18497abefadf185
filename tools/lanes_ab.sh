set -u
B="python bench.py --steps 20 --warmup 5 --no-ref-cuda --no-stress --no-fast-accept --no-cpu-baseline"
run() { # label, env...
  L=$1; shift
  env BENCH_NO_CLOCKS=1 "$@" timeout 40 $B > gpurun_out/s12_lanes_$L.json 2> gpurun_out/s12_lanes_$L.err
  python - $L <<'PY'
import json, sys
try:
    l = json.loads(open("gpurun_out/s12_lanes_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print("%-8s %.4f ms  e2e %.4f  launches %d  %s" % (sys.argv[1], l["ms_per_step"], l["e2e"]["ms_per_step"], l["gpu_launches"], {k: round(v, 3) for k, v in l["breakdown_ms"].items()}))
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/s12_lanes_%s.err" % sys.argv[1]).read()[-300:])
PY
}
run L2 RT_FRAME_LANES=2
run L2b RT_FRAME_LANES=2 RT_LANE_ORDER=1
run L3 RT_FRAME_LANES=3
run L4 RT_FRAME_LANES=4
run L4b RT_FRAME_LANES=4 RT_LANE_ORDER=1
run L8 RT_FRAME_LANES=8
timeout 60 python -m pytest tests/test_lanes.py -x -q 2>&1 | tail -8
