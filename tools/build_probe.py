#!/usr/bin/env python3
"""BVH build time of one scene (blocking rt_bvh_build, CUDA-event time inside the library), median of n, for the
one-launch cluster build (cluster sizes 16 and 8, with the per-phase times of CTA 0) and the multi-kernel pipeline.
  python tools/build_probe.py [scene] [grid_size]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["RT_BUILD_TIMERS"] = "1"
import rtb200_loader
pkg = rtb200_loader.load()
name = sys.argv[1] if len(sys.argv) > 1 else "world16"
grid = int(sys.argv[2]) if len(sys.argv) > 2 else 0
w = pkg.World(os.path.join(ROOT, "scenes", name + ".json"), grid_size=grid)
s = pkg.Scene(w)
PHASES = ["flatten+bounds", "morton", "sort", "karras", "local boxes", "spanning boxes", "collapse"]
for label, env in (("cluster16", {"RT_BUILD_CLUSTER": "16"}), ("cluster8", {"RT_BUILD_CLUSTER": "8"}), ("multi-kernel", {"RT_BUILD_PATH": "1"})):
    for k in ("RT_BUILD_CLUSTER", "RT_BUILD_PATH"):
        os.environ.pop(k, None)
    os.environ.update(env)
    for _ in range(5):
        s.build_bvh()
    ms = sorted(s.build_bvh()["ms_build"] for _ in range(41))
    bs = s.build_bvh()
    line = "%s grid %d %-12s: %d prims, %d nodes8, %d levels, %d launch(es), cluster %d, build median %.4f ms min %.4f" % (
        name, grid, label, bs["n_prims"], bs["n_bvh8_nodes"], bs["n_levels8"], bs["kernel_launches"],
        s.debug_get(pkg.RT_DEBUG_BUILD_CLUSTER), ms[20], ms[0])
    if s.debug_get(pkg.RT_DEBUG_BUILD_PATH) == 2:
        ph = [s.debug_get(pkg.RT_DEBUG_BUILD_PHASE_NS + i) / 1e3 for i in range(7)]
        line += " | phases us: " + ", ".join("%s %.1f" % (n, v) for n, v in zip(PHASES, ph))
        st = [s.debug_get(pkg.RT_DEBUG_BUILD_LEVEL_NS + i) for i in range(2 * bs["n_levels8"])]
        line += " | collapse levels us (A, B): " + " ".join("(%.1f, %.1f)" % ((st[2 * k] - (st[2 * k - 1] if k else st[0])) / 1e3, (st[2 * k + 1] - st[2 * k]) / 1e3) for k in range(bs["n_levels8"]))
    print(line, flush=True)
