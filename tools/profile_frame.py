#!/usr/bin/env python3
"""Renders a few frames of one workload through the C ABI (no torch): the command ncu wraps.
  python tools/profile_frame.py [scene] [width] [height] [depth] [frames] [flags] [grid_size]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader

pkg = rtb200_loader.load()
scene_name = sys.argv[1] if len(sys.argv) > 1 else "world16"
w = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
h = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
depth = int(sys.argv[4]) if len(sys.argv) > 4 else 9
frames = int(sys.argv[5]) if len(sys.argv) > 5 else 3
flags = int(sys.argv[6]) if len(sys.argv) > 6 else pkg.RT_RENDER_REBUILD_BVH
grid = int(sys.argv[7]) if len(sys.argv) > 7 else 0
world = pkg.World(os.path.join(ROOT, "scenes", scene_name + ".json"), width=w, height=h, depth=depth, grid_size=grid)
scene = pkg.Scene(world)
import ctypes
_rt = ctypes.CDLL('libcudart.so')
_p = ctypes.c_void_p()
assert _rt.cudaMalloc(ctypes.byref(_p), ctypes.c_size_t(4 * w * h)) == 0
dev_canvas = _p.value
for i in range(frames):
    canvas, st = scene.render(flags=flags | pkg.RT_RENDER_CANVAS_DEVICE, out_ptr=dev_canvas)
    print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in st.items()})
scene.close()
