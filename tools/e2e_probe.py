#!/usr/bin/env python3
"""Where an end-to-end banded frame spends its time (one GPU emulating rank 0 of `world`):
wall clock of the blocking calls vs the CUDA-event times inside the library."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader
pkg = rtb200_loader.load()
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
W, H = 3840, 2160
wf = pkg.World(os.path.join(ROOT, "scenes", "world16.json"), width=W, height=H, depth=9)
sc = pkg.Scene(wf)
desc = wf.desc
base = pkg.RT_RENDER_REBUILD_BVH
flags = base | pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_HOST_BANDS


def run(label, fn, n=30):
    for _ in range(5):
        fn()
    t = []; stats = []
    for _ in range(n):
        t0 = time.perf_counter(); st = fn(); t.append(1e3 * (time.perf_counter() - t0)); stats.append(st)
    t.sort()
    med = lambda k: sorted(s[k] for s in stats if s)[len(stats) // 2] if stats[0] else float("nan")
    print("%-58s wall med %.3f min %.3f | ms_total %.3f build %.3f trace %.3f" % (label, t[len(t) // 2], t[0], med("ms_total"), med("ms_build"), med("ms_trace")))


for shm in (None, "/rtb200_probe"):
    canvas, handle = pkg.host_canvas_open(W * H, shm_name=shm, create=True)
    run("bands rank0/%d -> %s canvas" % (world, "shm" if shm else "pinned"),
        lambda: sc.render(flags=flags, rank=0, world=world, tile_w=W, tile_h=4, out_ptr=canvas.ctypes.data)[1])
    pkg.host_canvas_close(handle, unlink=True)
import torch
slab = torch.zeros(sc.tiled_slab_pixels(sc.params(flags=flags, world=world, tile_w=W, tile_h=4)), dtype=torch.int32, device="cuda")
run("bands rank0/%d -> device slab" % world,
    lambda: sc.render(flags=base | pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_CANVAS_DEVICE, rank=0, world=world, tile_w=W, tile_h=4, out_ptr=slab.data_ptr())[1])
run("same, 32x16 tiles", lambda: sc.render(flags=base | pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_CANVAS_DEVICE, rank=0, world=world, tile_w=32, tile_h=16, out_ptr=slab.data_ptr())[1])
run("same, BVH kept", lambda: sc.render(flags=pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_CANVAS_DEVICE, rank=0, world=world, tile_w=W, tile_h=4, out_ptr=slab.data_ptr())[1])
run("set_instances + set_camera", lambda: (sc.set_instances(desc.instances, desc.n_instances), sc.set_camera(desc.camera), None)[2])
run("set_instances + refit frame", lambda: (sc.set_instances(desc.instances, desc.n_instances), sc.render(flags=pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_CANVAS_DEVICE, rank=0, world=world, tile_w=W, tile_h=4, out_ptr=slab.data_ptr())[1])[1])
sc.build_bvh()
t0 = time.perf_counter()
for _ in range(20):
    bs = sc.build_bvh()
print("rt_bvh_build blocking: wall %.3f ms, ms_build %.3f, launches %d" % (1e3 * (time.perf_counter() - t0) / 20, bs["ms_build"], bs["kernel_launches"]))
