timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
CMD="python tools/profile_frame.py world8_stress 3840 2160 2 2"
timeout 120 $CMD > gpurun_out/plain_tmp.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_resolve|k_path_trace|k_path_advance|k_shadow" -s 4 -c 8 -o gpurun_out/prof_r1d_stress $CMD > gpurun_out/ncu_full_d.log 2>&1
tail -2 gpurun_out/ncu_full_d.log
ncu -i gpurun_out/prof_r1d_stress.ncu-rep --page details --csv > gpurun_out/prof_r1d_stress_details.csv 2>/dev/null
ls -la gpurun_out/
