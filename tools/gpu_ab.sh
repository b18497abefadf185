# A/B of prebuilt engine variants: tools/gpu_ab.sh variant_X.so ...
cd gpu-ray-tracer_b200; cp librtb200.so variant_A.so; cd ..
for v in variant_A.so "$@"; do
  cp gpu-ray-tracer_b200/$v gpu-ray-tracer_b200/librtb200.so
  echo "== $v"
  python tools/profile_frame.py world16 3840 2160 9 6 | tail -2 | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print({k:d[k] for k in ('ms_build','ms_trace','ms_primary','ms_shade','ms_shadow','ms_bounce')})"
  python tools/profile_frame.py world8_stress 3840 2160 2 6 | tail -2 | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print({k:d[k] for k in ('ms_build','ms_trace','ms_primary','ms_shade','ms_shadow','ms_bounce')})"
done
cp gpu-ray-tracer_b200/variant_A.so gpu-ray-tracer_b200/librtb200.so
