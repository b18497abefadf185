#!/usr/bin/env python3
"""Parity and speed of the opt-in tolerance mode (RT_RENDER_FAST_ACCEPT) against the bit-exact default, on BASELINE.json's five
configurations: closest-hit id agreement on the primary rays, largest relative error of the hit distance, image RMSE on [0, 1]
colour (north_star's three tolerances: >= 99.99 %, 1e-4, 1e-3), exact tests per ray and frame time (device canvas, BVH rebuilt,
median of 15 frames). The default itself equals the CPU oracle bit for bit (tests/), so it stands in for the reference here.
  python tools/fast_accept_parity.py > profiles/r2_fast_accept_parity.jsonl"""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader

pkg = rtb200_loader.load()
CASES = [("C1", "world1", dict(width=640, height=480, depth=1)), ("C2", "world4", dict(width=1920, height=1080, depth=5)),
         ("C3", "world8", dict(width=3840, height=2160)), ("C4", "world16", dict(width=3840, height=2160, depth=9)),
         ("C5", "world8_stress", dict(width=3840, height=2160, depth=9)), ("f3", "grid64_mirror", dict(width=1920, height=1080))]
rt = ctypes.CDLL("libcudart.so")


def channels(a):
    return np.stack([(a >> s) & 255 for s in (24, 16, 8, 0)], -1).astype(np.float64) / 255.0


for tag, name, kw in CASES:
    w = pkg.World(os.path.join(ROOT, "scenes", name + ".json"), **kw)
    s = pkg.Scene(w)
    W, H = w.width, w.height
    dev = ctypes.c_void_p()
    assert rt.cudaMalloc(ctypes.byref(dev), ctypes.c_size_t(4 * W * H)) == 0
    out = {}
    for mode, flag in (("exact", 0), ("fast", pkg.RT_RENDER_FAST_ACCEPT)):
        os.environ["RT_FAST_ACCEPT"] = "1" if flag else "0"
        canvas, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_STATS | flag)
        inst, tri, t = s.trace_primary(use_bvh=1)
        ms = []
        for _ in range(15):
            _, st2 = s.render(flags=pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE | flag, out_ptr=dev.value)
            ms.append(st2["ms_total"])
        ms.sort()
        out[mode] = dict(canvas=canvas, inst=inst, tri=tri, t=t, st=st, ms=ms[len(ms) // 2], st2=st2)
    os.environ["RT_FAST_ACCEPT"] = "0"
    e, f = out["exact"], out["fast"]
    same = (e["inst"] == f["inst"]) & (e["tri"] == f["tri"])
    hit = same & (e["inst"] >= 0)
    rel = np.abs(f["t"][hit].astype(np.float64) - e["t"][hit]) / np.abs(e["t"][hit])
    line = {"config": tag, "scene": name, **kw, "width": W, "height": H, "pixels": int(W * H),
            "ids_identical_pct": 100.0 * float(same.mean()), "ids_differ": int((~same).sum()),
            "t_max_rel_err": float(rel.max()) if rel.size else 0.0, "t_bit_identical_pct": 100.0 * float((f["t"][hit] == e["t"][hit]).mean()) if rel.size else 100.0,
            "rmse": float(np.sqrt(np.mean((channels(e["canvas"]) - channels(f["canvas"])) ** 2))),
            "pixels_differ": int((e["canvas"] != f["canvas"]).sum()),
            "rays": e["st"]["rays_total"], "rays_fast": f["st"]["rays_total"],
            "exact_tests_per_ray": e["st"]["tris_tested"] / e["st"]["rays_total"], "exact_tests_per_ray_fast": f["st"]["tris_tested"] / f["st"]["rays_total"],
            "ms_exact": e["ms"], "ms_fast": f["ms"], "speedup": e["ms"] / f["ms"],
            "ms_primary": [e["st2"]["ms_primary"], f["st2"]["ms_primary"]], "ms_shadow": [e["st2"]["ms_shadow"], f["st2"]["ms_shadow"]],
            "within_north_star": bool(same.mean() >= 0.9999 and (rel.max() if rel.size else 0) <= 1e-4)}
    line["within_north_star"] = bool(line["within_north_star"] and line["rmse"] <= 1e-3)
    print(json.dumps(line), flush=True)
    rt.cudaFree(dev)
    s.close()
