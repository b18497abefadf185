CMD="python tools/profile_frame.py world8_stress 3840 2160 2 3"
timeout 120 $CMD > gpurun_out/plain_tmp.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_stress.csv $CMD > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/launches_stress.csv
