# session 14, call F: launch lists at HEAD (world16 4K and world16 at grid 256), ncu --set full of the frame kernels on the big world
mkdir -p gpurun_out
for cfg in "world16 0 w16" "world16 256 g256"; do set -- $cfg
  CMD="python tools/profile_frame.py $1 3840 2160 9 3 1 $2"
  timeout 200 $CMD > gpurun_out/plain_s14f_$3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_s14f_$3.csv $CMD > /dev/null 2>&1
  echo "== $3"; tail -1 gpurun_out/plain_s14f_$3.log | cut -c1-400; python tools/launch_summary.py gpurun_out/launches_s14f_$3.csv
done
CMD="python tools/profile_frame.py world16 3840 2160 9 2 1 256"
ncu --set full --clock-control none --import-source on -k regex:"k_primary|k_shadow|k_resolve|k_collapse|k_radix_scatter" -s 6 -c 12 -o gpurun_out/prof_s14f_g256 $CMD > gpurun_out/ncu_full_s14f.log 2>&1
tail -2 gpurun_out/ncu_full_s14f.log
ncu -i gpurun_out/prof_s14f_g256.ncu-rep --page details --csv > gpurun_out/prof_s14f_g256_details.csv 2>/dev/null
ls -la gpurun_out/prof_s14f_g256*
