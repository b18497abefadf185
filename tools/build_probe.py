#!/usr/bin/env python3
"""BVH build time of one scene (blocking rt_bvh_build, CUDA-event time inside the library), median of n."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader
pkg = rtb200_loader.load()
name = sys.argv[1] if len(sys.argv) > 1 else "world16"
grid = int(sys.argv[2]) if len(sys.argv) > 2 else 0
w = pkg.World(os.path.join(ROOT, "scenes", name + ".json"), grid_size=grid)
s = pkg.Scene(w)
for _ in range(5):
    s.build_bvh()
ms = sorted(s.build_bvh()["ms_build"] for _ in range(41))
bs = s.build_bvh()
print("%s grid %d: %d prims, %d nodes8, %d launches, build median %.4f ms min %.4f (RT_COLLAPSE_BLOCKS=%s)" % (
    name, grid, bs["n_prims"], bs["n_bvh8_nodes"], bs["kernel_launches"], ms[20], ms[0], os.environ.get("RT_COLLAPSE_BLOCKS", "-")))
