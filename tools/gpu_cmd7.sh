CMD="python tools/profile_frame.py world8_stress 3840 2160 2 2"
timeout 120 $CMD > gpurun_out/plain_tmp.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_resolve|k_path_advance|k_path_trace" -s 3 -c 3 -o gpurun_out/prof_r1e_stress $CMD > gpurun_out/ncu_full_e.log 2>&1
tail -2 gpurun_out/ncu_full_e.log
ncu -i gpurun_out/prof_r1e_stress.ncu-rep --page details --csv > gpurun_out/prof_r1e_stress_details.csv 2>/dev/null
ncu -i gpurun_out/prof_r1e_stress.ncu-rep --page source --csv --kernel-name regex:k_resolve > gpurun_out/prof_r1e_resolve_source.csv 2>/dev/null
