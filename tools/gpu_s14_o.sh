# session 14, call O: launch list + ncu --set full at HEAD (world16 4K depth 9), for profiles/
mkdir -p gpurun_out
CMD="python tools/profile_frame.py world16 3840 2160 9 3 1"
timeout 200 $CMD > gpurun_out/plain_s14o.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_s14o_w16.csv $CMD > /dev/null 2>&1
tail -1 gpurun_out/plain_s14o.log | cut -c1-300; python tools/launch_summary.py gpurun_out/launches_s14o_w16.csv
ncu --set full --clock-control none --import-source on -k regex:"k_primary|k_shadow|k_resolve" -s 4 -c 4 -o gpurun_out/prof_s14o_w16 $CMD > gpurun_out/ncu_full_s14o.log 2>&1
tail -1 gpurun_out/ncu_full_s14o.log
ncu -i gpurun_out/prof_s14o_w16.ncu-rep --page details --csv > gpurun_out/prof_s14o_w16_details.csv 2>/dev/null
ncu -i gpurun_out/prof_s14o_w16.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,sm__inst_executed.sum > gpurun_out/prof_s14o_w16_raw.csv 2>/dev/null
cat gpurun_out/prof_s14o_w16_raw.csv | cut -c1-400 | tail -6
