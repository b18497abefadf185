CMD="python tools/profile_frame.py world16 3840 2160 9 3"
timeout 120 $CMD > gpurun_out/plain_r1c.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_r1c.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_shade|k_primary" -s 2 -c 2 -o gpurun_out/prof_r1c $CMD > gpurun_out/ncu_full_c.log 2>&1
tail -2 gpurun_out/ncu_full_c.log
