#!/usr/bin/env python3
"""The reference's own CUDA build (oracle/_ref/ref_gpu: unmodified sources, nvcc -arch=sm_100) beside this engine on the
same GPU and the same scene files: frame time in the reference's bench() scope (BVH build + trace), Mrays/s (rays
counted by this engine's frame of the same scene: both answer the same closest-hit queries) and the pixel difference
between the two canvases. Shipped resolution for four scenes, then BASELINE.json's 4K configurations.
  python tools/ref_gpu_compare.py [frames] > profiles/r1_ref_gpu_compare.jsonl"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader

pkg = rtb200_loader.load()
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 20
ref_gpu = os.path.join(ROOT, "oracle", "_ref", "ref_gpu")
CASES = [("world4", {}), ("world8", {}), ("world16", {}), ("world8_stress", {}),
         ("world16", {"width": 3840, "height": 2160, "depth": 9}), ("world8_stress", {"width": 3840, "height": 2160, "depth": 9})]
for name, override in CASES:
    path = os.path.join(ROOT, "scenes", name + ".json")
    tmp_json = None
    if override:
        cfg = json.load(open(path))
        cfg.update(override)
        with tempfile.NamedTemporaryFile("w", suffix=".json", delete=False) as tf:
            json.dump(cfg, tf)
            tmp_json = path = tf.name
    with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
        dump = f.name
    try:
        out = subprocess.run([ref_gpu, path, "--gpu", "--frames", str(frames), "--kd", "16", "--canvas", dump],
                             capture_output=True, text=True, timeout=600)
        info = json.loads(out.stdout.strip().splitlines()[-1])
        raw = np.fromfile(dump, dtype=np.uint32)
        w, h = int(raw[0]), int(raw[1])
        ref_canvas = raw[2:].reshape(h, w)
        world = pkg.World(path)
    finally:
        os.unlink(dump)
        if tmp_json:
            os.unlink(tmp_json)
    scene = pkg.Scene(world)
    ms = []
    for _ in range(frames):
        canvas, st = scene.render(flags=pkg.RT_RENDER_REBUILD_BVH)
        ms.append(st["ms_total"])
    ms.sort()
    ch = lambda a: np.stack([(a >> s) & 255 for s in (24, 16, 8, 0)], -1).astype(np.float64) / 255.0
    diff = canvas != ref_canvas
    line = {"scene": name, "override": override, "width": w, "height": h, "rays_per_frame": st["rays_total"],
            "ref_gpu_ms_median": info["ms_median"], "ref_gpu_ms_min": info["ms_min"],
            "ref_gpu_mrays_s": st["rays_total"] / info["ms_median"] / 1e3,
            "ours_ms_median": ms[len(ms) // 2], "ours_mrays_s": st["rays_total"] / ms[len(ms) // 2] / 1e3,
            "pixels_differ": int(diff.sum()), "pixels": int(diff.size),
            "max_channel_diff_255": int(np.abs(ch(canvas) - ch(ref_canvas)).max() * 255 + 0.5),
            "rmse": float(np.sqrt(np.mean((ch(canvas) - ch(ref_canvas)) ** 2))),
            "note": "ours: rt_render into a pageable host canvas, BVH rebuilt, CUDA events over build + trace + copy; "
                    "ref_gpu: std::chrono around rtracer::gpu::update_scene (its bench() scope), -d 16"}
    print(json.dumps(line), flush=True)
    scene.close()
