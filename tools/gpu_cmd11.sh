timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for sc in "world8_stress 3840 2160 9" "world8_stress 640 480 9"; do
python tools/profile_frame.py $sc 4 1 2>&1 | tail -1 | cut -c1-900
python tools/profile_frame.py $sc 2 33 2>&1 | tail -1 | cut -c400-900
done
