# A/B of prebuilt engine variants: tools/gpu_ab.sh variant_X.so ...
cd gpu-ray-tracer_b200; cp librtb200.so variant_A.so; cd ..
for v in variant_A.so "$@"; do
  cp gpu-ray-tracer_b200/$v gpu-ray-tracer_b200/librtb200.so
  echo "== $v"
  for sc in "world8_stress 3840 2160 2" "world8_stress 640 480 2"; do
  python tools/profile_frame.py $sc 6 | tail -2 | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print({k:d[k] for k in ('ms_build','ms_trace','ms_primary','ms_shade','ms_shadow','ms_bounce')})"
  done
done
cp gpu-ray-tracer_b200/variant_A.so gpu-ray-tracer_b200/librtb200.so
