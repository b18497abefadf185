#!/usr/bin/env python3
"""Golden fixtures for per-frame motion, produced by RUNNING THE UNMODIFIED REFERENCE:
oracle/_ref/ref_cpu --moves applies camera key presses, mouse look and instance translate / rotate through
the reference's own Entity API (arithmetic of src/main.cc:140-177, include/rayprimitives/entity.h:49-76) and
then traces one frame with rtracer::cpu::update_scene. Only runs in the build container.

Per case: moved_<name>.json — the move list, sha256 of the moved scene dump, canvas hash
          moved_<name>.npz  — RGBA8 canvas (serial semantics) and per-pixel primary closest hit with the leaf-box rule."""
import hashlib
import json
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
REF_CPU = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")

CASES = {
    # camera walk + look, two translated cubes, two cubes rotated in place (the shipped worlds only have identity
    # rotations, so this is the one pinned case of the general quaternion path), one cube lifted off the terrain
    "world8": "w,w,d,m:3:-2,m:1:4,s,g:5:0.25:0.5:-0.125,l:17:0.1:0:0.2,r:40:0:1:0:0.3,r:41:1:0:0:0.2,g:100:0:1:0",
    "world8_stress": "a,a,w,m:-5:1,m:2:2,g:300:0:0.75:0,r:301:0:0:1:0.5,l:302:0.3:0.1:0",
    "world4": "m:0:7,s,s,d,r:3:0.6:0.8:0:1.1,r:3:0:0:1:0.4,g:10:-0.5:0:0.5",
}


def run(name, moves):
    with tempfile.TemporaryDirectory() as td:
        cpath, ipath, spath = (os.path.join(td, f) for f in ("c.bin", "i.bin", "s.txt"))
        cmd = [REF_CPU, name + ".json", "--moves", moves, "--canvas", cpath, "--ids", ipath, "--ids-bvh-only", "--scene", spath]
        out = subprocess.run(cmd, cwd=REF, capture_output=True, text=True, check=True).stdout
        info = json.loads(out.strip().splitlines()[-1])
        raw = np.fromfile(cpath, dtype=np.int32)
        w, h = int(raw[0]), int(raw[1])
        canvas = raw[2:].view(np.uint32).reshape(h, w)
        ids = np.fromfile(ipath, dtype=np.uint8)[8:]
        rec = np.frombuffer(ids.tobytes(), dtype=np.dtype([("trans", "<i4"), ("tri", "<i4"), ("t", "<u4"),
                                                            ("trans_bf", "<i4"), ("tri_bf", "<i4"), ("t_bf", "<u4")]))
        np.savez_compressed(os.path.join(HERE, "moved_" + name + ".npz"), canvas=canvas, inst=rec["trans"].astype(np.int16),
                            tri=rec["tri"].astype(np.int8), t_bits=rec["t"].copy())
        dump = open(spath, "rb").read()
        meta = {"scene": name + ".json", "moves": moves, "scene_sha256": hashlib.sha256(dump).hexdigest(),
                "scene_bytes": len(dump), "hash": info["hash"], "nonzero": info["nonzero"], "width": w, "height": h,
                "semantics": "serial (rtracer::cpu::update_scene)",
                "generator": "tests/golden/make_moved.py via oracle/_ref/ref_cpu --moves (unmodified reference sources)"}
        json.dump(meta, open(os.path.join(HERE, "moved_" + name + ".json"), "w"), indent=1)
        print(name, meta["hash"], meta["nonzero"])


if __name__ == "__main__":
    for n, m in CASES.items():
        run(n, m)
