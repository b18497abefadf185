#!/usr/bin/env python3
"""Longest single work unit of the two traversal kernels (stats variants) beside the kernels' durations."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import rtb200_loader
pkg = rtb200_loader.load()
rt = ctypes.CDLL("libcudart.so")
dev = ctypes.c_void_p(); assert rt.cudaMalloc(ctypes.byref(dev), ctypes.c_size_t(4 * 3840 * 2160)) == 0
for (W, H) in ((3840, 2160), (1920, 1080)):
    w = pkg.World(os.path.join(ROOT, "scenes", "world16.json"), width=W, height=H, depth=9)
    s = pkg.Scene(w)
    for i in range(6):
        _, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE, out_ptr=dev.value)
    _, ss = s.render(flags=pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE | pkg.RT_RENDER_STATS, out_ptr=dev.value)
    v = [round(x * 1e6) for x in ss["ms_wave_phase"][:3]]
    print("   worst shadow unit: queue position %d of %d, %d node fetches, most exact tests of a lane %d, most triangle filters of a lane %d" % (
        v[0], 2 * ((ss["n_hits"] + 31) // 32 * 32), v[1] // 10000, v[1] % 10000, v[2]))
    print("%dx%d: k_tile<1> %.3f ms (stats build %.3f), longest tile %.3f ms | k_shadow %.3f ms (stats %.3f), longest unit %.3f ms" % (
        W, H, st["ms_primary"], ss["ms_primary"], ss["ms_wave_phase"][2], st["ms_shadow"], ss["ms_shadow"], ss["ms_wave_phase"][3]))
    s.close()
