# HEAD check on one B200: GPU tests, smoke, bench (both headline scenes), launch list, ncu --set full of the frame's kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r1h.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_r1h.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 50 --warmup 5 2>gpurun_out/bench_r1h_n1.err | tail -1 > gpurun_out/bench_r1h_n1.json; cut -c1-1500 gpurun_out/bench_r1h_n1.json
timeout 600 python bench.py --steps 50 --warmup 5 --scene world8_stress --no-cpu-baseline 2>gpurun_out/bench_r1h_stress.err | tail -1 > gpurun_out/bench_r1h_stress_n1.json; cut -c1-600 gpurun_out/bench_r1h_stress_n1.json
CMD="python tools/profile_frame.py world16 3840 2160 9 3"
timeout 120 $CMD > gpurun_out/plain_r1h.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_r1h_world16.csv $CMD > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/launches_r1h_world16.csv
ncu --set full --clock-control none --import-source on -k regex:"k_primary|k_shadow|k_resolve" -s 4 -c 4 -o gpurun_out/prof_r1h_world16 $CMD > gpurun_out/ncu_full_r1h.log 2>&1
tail -2 gpurun_out/ncu_full_r1h.log
ncu -i gpurun_out/prof_r1h_world16.ncu-rep --page details --csv > gpurun_out/prof_r1h_world16_details.csv 2>/dev/null
CMD2="python tools/profile_frame.py world8_stress 3840 2160 9 3"
timeout 120 $CMD2 > gpurun_out/plain_r1h_stress.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_r1h_stress.csv $CMD2 > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/launches_r1h_stress.csv
