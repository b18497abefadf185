CMD="python tools/profile_frame.py world16 3840 2160 9 3"
timeout 120 $CMD > gpurun_out/plain_tmp.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_tmp.csv $CMD > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/launches_tmp.csv
