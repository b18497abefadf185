timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for sc in "world16 3840 2160 9" "world8_stress 3840 2160 2" "world8_stress 640 480 2" "world1 640 480 2"; do
python tools/profile_frame.py $sc 6 | tail -2 | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print({k:d[k] for k in ('ms_build','ms_trace','ms_primary','ms_shade','ms_shadow','ms_bounce','waves','kernel_launches')})"
done
