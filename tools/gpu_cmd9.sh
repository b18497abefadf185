for sc in "world8_stress 3840 2160 2" "world8_stress 640 480 2"; do
RTB200_DEBUG_STATS=1 python tools/profile_frame.py $sc 2 33 2>&1 | tail -2 | cut -c1-700
done
