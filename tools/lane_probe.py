#!/usr/bin/env python3
"""Concurrent lanes (rt_scene_set_lanes) on ONE rank's share of a sharded frame, on one GPU — the per-rank frame time of an
N-GPU run without the box (tools/shard_probe.py showed that a rank's tiles cost the same alone on one GPU as on the N-GPU
box): python tools/lane_probe.py [world ...]   ->   median ms of trace (all frame kernels) and of the whole step per lane count."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader
pkg = rtb200_loader.load()
worlds = [int(x) for x in sys.argv[1:]] or [8, 4, 2, 1]
w = pkg.World(os.path.join(ROOT, "scenes", "world16.json"), width=3840, height=2160, depth=9)
s = pkg.Scene(w)
rt = ctypes.CDLL("libcudart.so")
dev = ctypes.c_void_p()
assert rt.cudaMalloc(ctypes.byref(dev), ctypes.c_size_t(4 * 3840 * 2160)) == 0
med = lambda xs: sorted(xs)[len(xs) // 2]
for world in worlds:
    flags = pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE | (pkg.RT_RENDER_TILED_OUTPUT if world > 1 else 0)
    for rank in sorted({0, world - 1}):
        kw = dict(flags=flags, rank=rank, world=world, tile_w=32, tile_h=16, out_ptr=dev.value)
        s.set_lanes(1)
        for _ in range(8):
            s.render(**kw)  # the shadow-kernel tuner of this tile shape settles
        res = {}
        for rep in range(3):
            for lanes in (1, 2, 3, 4):
                s.set_lanes(lanes)
                for _ in range(5):
                    _, st = s.render(**kw)
                    res.setdefault(lanes, []).append((st["ms_trace"], st["ms_total"]))
                assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == lanes
        print("world %d rank %d: " % (world, rank) + "  ".join(
            "L%d trace %.4f step %.4f" % (l, med([a for a, _ in v]), med([b for _, b in v])) for l, v in sorted(res.items())), flush=True)
