#!/usr/bin/env python3
"""Kernel times of ONE rank's share of a sharded frame, on one GPU: python tools/shard_probe.py [world] [tile_w] [tile_h]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader
pkg = rtb200_loader.load()
world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
tw = int(sys.argv[2]) if len(sys.argv) > 2 else 32
th = int(sys.argv[3]) if len(sys.argv) > 3 else 16
w = pkg.World(os.path.join(ROOT, "scenes", "world16.json"), width=3840, height=2160, depth=9)
s = pkg.Scene(w)
rt = ctypes.CDLL("libcudart.so")
dev = ctypes.c_void_p()
assert rt.cudaMalloc(ctypes.byref(dev), ctypes.c_size_t(4 * 3840 * 2160)) == 0
flags = pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE | (pkg.RT_RENDER_TILED_OUTPUT if world > 1 else 0)
for rank in range(min(world, 2)):
    ms = []
    for i in range(12):
        _, st = s.render(flags=flags, rank=rank, world=world, tile_w=tw, tile_h=th, out_ptr=dev.value)
        ms.append(st)
    st = ms[-1]
    f = lambda k: sorted(x[k] for x in ms[6:])[len(ms[6:]) // 2]
    print("world %d rank %d tile %dx%d: hits %d shadow rays %d culled %d | k_tile<1> %.3f  shade %.3f  k_shadow %.3f  total %.3f ms  (BEAM_SHADOW=%s)" % (
        world, rank, tw, th, st["n_hits"], st["rays_shadow"], st["rays_culled"], f("ms_primary"), f("ms_shade"), f("ms_shadow"), f("ms_total"), os.environ.get("RT_BEAM_SHADOW", "auto")))
