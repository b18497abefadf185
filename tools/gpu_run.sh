#!/bin/bash
# One parameterised GPU-box script (replaces the per-session tools/gpu_*.sh of round 1). Run under gpurun from the repo
# root:   gpurun --timeout 1500 -- 'bash tools/gpu_run.sh <tag> <step> [<step> ...]'
# Every step writes into gpurun_out/<tag>_*; copy what should be judged into profiles/ afterwards.
#   tests           python -m pytest tests -m gpu -x -q
#   tests:<expr>    ... -k <expr>
#   bench           python bench.py --steps 20 --warmup 5                 (world16 4K depth 9, N = 1)
#   bench:<args>    python bench.py <args>   (commas become spaces, e.g. bench:--scene,world8_stress)
#   multi:<N>       torchrun --nproc-per-node N bench.py --gpus N
#   launches        ncu launch list (gpu__time_duration.sum) of tools/profile_frame.py world16 3840 2160 9 3
#   ncu:<regex>     ncu --set full of the kernels matching <regex> in the same command (summary csv + .ncu-rep)
#   sanitizer       compute-sanitizer memcheck over the smoke frame and a small cluster build
#   refcmp          tools/ref_gpu_compare.py (the reference's own CUDA build beside the engine)
#   cmd:<shell>     anything else (commas become spaces)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=$1; shift
export PYTHONUNBUFFERED=1
for STEP in "$@"; do
  NAME=${STEP%%:*}; ARG=""; [[ "$STEP" == *:* ]] && ARG=${STEP#*:}
  ARGS=${ARG//,/ }
  echo "=== $TAG $STEP"
  case $NAME in
    tests)
      if [ -n "$ARG" ]; then timeout 1200 python -m pytest tests -m gpu -x -q -k "$ARGS" 2>&1 | tail -40 | tee gpurun_out/${TAG}_tests.log
      else timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 | tee gpurun_out/${TAG}_tests.log; fi ;;
    bench)
      SUF=$(echo "$ARG" | tr -c 'A-Za-z0-9' '_' | cut -c1-40)
      timeout 600 python bench.py ${ARGS:---steps 20 --warmup 5} > gpurun_out/${TAG}_bench_${SUF}.json 2> gpurun_out/${TAG}_bench_${SUF}.err
      tail -c 1500 gpurun_out/${TAG}_bench_${SUF}.json; tail -5 gpurun_out/${TAG}_bench_${SUF}.err ;;
    multi)
      N=${ARG%%,*}; EXTRA=""; [[ "$ARG" == *,* ]] && EXTRA=${ARG#*,}; EXTRA=${EXTRA//,/ }
      SUF=$(echo "$EXTRA" | tr -c 'A-Za-z0-9' '_' | cut -c1-30)
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
        bench.py --gpus $N --steps 20 --warmup 5 $EXTRA > gpurun_out/${TAG}_bench_n${N}_${SUF}.json 2> gpurun_out/${TAG}_bench_n${N}_${SUF}.err
      tail -c 1200 gpurun_out/${TAG}_bench_n${N}_${SUF}.json; tail -5 gpurun_out/${TAG}_bench_n${N}_${SUF}.err ;;
    multiab)
      # multiab:<N>;<label>=<env and bench args, comma separated>;...   several N-GPU bench runs in one call, e.g.
      #   multiab:8;tile=--gather,fused-root;split=--split-shade,--gather,p2p
      N=${ARG%%;*}; REST=${ARG#*;}
      IFS=';' read -ra RUNS <<< "$REST"
      for RUN in "${RUNS[@]}"; do
        LBL=${RUN%%=*}; RA=${RUN#*=}; RA=${RA//,/ }
        ENVS=""; BA=""
        for TOK in $RA; do if [[ "$TOK" == [A-Z_]*=* && "$TOK" != --* ]]; then ENVS="$ENVS $TOK"; else BA="$BA $TOK"; fi; done
        env $ENVS timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
          bench.py --gpus $N --steps 20 --warmup 5 $BA > gpurun_out/${TAG}_n${N}_${LBL}.json 2> gpurun_out/${TAG}_n${N}_${LBL}.err
        python - "$LBL" gpurun_out/${TAG}_n${N}_${LBL}.json <<'PY'
import json, sys
try:
    l = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print("%-14s N=%d %.3f ms/step  %.0f Mrays/s  e2e %.3f ms  launches %d  %s" % (sys.argv[1], l["n_gpus"], l["ms_per_step"], l["value"], l["e2e"]["ms_per_step"], l["gpu_launches"], {k: round(v, 3) for k, v in l["breakdown_ms"].items()}))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
        grep -v "^W\|^\*\*\*\|OMP_NUM" gpurun_out/${TAG}_n${N}_${LBL}.err | tail -4
      done ;;
    launches)
      timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv \
        python tools/profile_frame.py ${ARGS:-world16 3840 2160 9 3 1} > gpurun_out/${TAG}_launches.log 2>&1
      python tools/launch_summary.py gpurun_out/${TAG}_launches.csv | tee gpurun_out/${TAG}_launches.txt ;;
    ncu)
      # ncu:<regex>[,<skip>,<count>[,<profile_frame args...>]]   (skip / count are launches MATCHING the regex)
      RX=${ARG%%,*}; REST=""; [[ "$ARG" == *,* ]] && REST=${ARG#*,}
      SKIP=1; CNT=2; PF="world16 3840 2160 9 3 1"
      if [ -n "$REST" ]; then SKIP=${REST%%,*}; REST2=${REST#*,}; CNT=${REST2%%,*}; [[ "$REST2" == *,* ]] && PF=${REST2#*,} && PF=${PF//,/ }; fi
      timeout 900 env ${NCU_ENV:-} ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c $CNT -f -o gpurun_out/${TAG}_full \
        python tools/profile_frame.py $PF > gpurun_out/${TAG}_ncu.log 2>&1
      ncu -i gpurun_out/${TAG}_full.ncu-rep --page details --csv > gpurun_out/${TAG}_ncu_details.csv 2>/dev/null
      ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_ncu_raw.csv 2>/dev/null
      ls -la gpurun_out/${TAG}_full.ncu-rep ;;
    ncum)
      # ncum[:<env>,...]  — byte counters of every kernel of one frame: DRAM, L2 (lts) and L1 (l1tex) traffic, duration, instructions
      timeout 900 env ${ARGS} ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct \
        --clock-control none -s 5 -c 12 --csv --log-file gpurun_out/${TAG}_ncum.csv python tools/profile_frame.py world16 3840 2160 9 3 1 > gpurun_out/${TAG}_ncum.log 2>&1
      tail -2 gpurun_out/${TAG}_ncum.log ;;
    sanitizer)
      timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -25 | tee gpurun_out/${TAG}_sanitizer.log ;;
    refcmp)
      timeout 900 python tools/ref_gpu_compare.py ${ARGS:-20} > gpurun_out/${TAG}_ref_gpu_compare.jsonl 2> gpurun_out/${TAG}_ref_gpu_compare.err
      cat gpurun_out/${TAG}_ref_gpu_compare.jsonl | cut -c1-600 ;;
    ab)
      # ab:<variant>[,<variant>...][;bench args]  — A/B of engine builds from tools/build_variant.sh ("main" = the shipped library)
      VARS=${ARG%%;*}; BARGS=""; [[ "$ARG" == *";"* ]] && BARGS=${ARG#*;}; BARGS=${BARGS//,/ }
      for V in ${VARS//,/ }; do
        LIB=""; [ "$V" != "main" ] && LIB=gpu-ray-tracer_b200/variants/librtb200_${V}.so
        RTB200_CUDA_LIB=$LIB timeout 600 python bench.py --steps 20 --warmup 5 --no-stress --no-ref-cuda --no-cpu-baseline --no-fast-accept $BARGS \
          > gpurun_out/${TAG}_ab_${V}.json 2> gpurun_out/${TAG}_ab_${V}.err
        python - "$V" gpurun_out/${TAG}_ab_${V}.json <<'PY'
import json, sys
try:
    l = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print("%-12s %.3f ms/step  %.0f Mrays/s  e2e %.3f ms  %s" % (sys.argv[1], l["ms_per_step"], l["value"], l["e2e"]["ms_per_step"], {k: round(v, 3) for k, v in l["breakdown_ms"].items()}))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
        tail -3 gpurun_out/${TAG}_ab_${V}.err
      done ;;
    abf)
      # abf:<label>=<env / bench args, comma separated>;...   A/B of bench flags on one GPU with the shipped library
      IFS=';' read -ra RUNS <<< "$ARG"
      for RUN in "${RUNS[@]}"; do
        LBL=${RUN%%=*}; RA=${RUN#*=}; [ "$RA" == "$RUN" ] && RA=""; RA=${RA//,/ }
        ENVS=""; BA=""
        for TOK in $RA; do if [[ "$TOK" == [A-Z_]*=* && "$TOK" != --* ]]; then ENVS="$ENVS $TOK"; else BA="$BA $TOK"; fi; done
        env $ENVS timeout 600 python bench.py --steps 20 --warmup 5 --no-stress --no-ref-cuda --no-cpu-baseline --no-fast-accept $BA \
          > gpurun_out/${TAG}_abf_${LBL}.json 2> gpurun_out/${TAG}_abf_${LBL}.err
        python - "$LBL" gpurun_out/${TAG}_abf_${LBL}.json <<'PY'
import json, sys
try:
    l = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print("%-14s %.3f ms/step  %.0f Mrays/s  e2e %.3f ms  launches %d  %s" % (sys.argv[1], l["ms_per_step"], l["value"], l["e2e"]["ms_per_step"], l["gpu_launches"], {k: round(v, 3) for k, v in l["breakdown_ms"].items()}))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
        tail -2 gpurun_out/${TAG}_abf_${LBL}.err
      done ;;
    cmd)
      timeout 1200 bash -c "$ARGS" 2>&1 | tail -60 | tee gpurun_out/${TAG}_cmd.log ;;
    *) echo "unknown step $STEP" ;;
  esac
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,temperature.gpu --format=csv,noheader | head -8
