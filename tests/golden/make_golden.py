#!/usr/bin/env python3
"""Generates the golden fixtures in tests/golden/ by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref/ref_cpu, built from /root/reference by oracle/ref_build/Makefile). Only runs in the
build container (needs /root/reference); the fixtures it writes are committed and travel.

Per world: <name>.json  — canvas hash / non-zero pixel count / instance count / frame time
           <name>.npz   — RGBA8 canvas (serial-path semantics), per-pixel primary-ray closest hit
                          (instance, triangle, t) with the BVH leaf-box filter, and (small worlds)
                          brute force.
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
REF_CPU = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")

WORLDS = {  # name -> (json, brute-force ids too?)
    "world1": ("world1.json", True), "world2": ("world2.json", True), "world4": ("world4.json", True),
    "world8": ("world8.json", True), "world16": ("world16.json", False),
    "world8_stress": ("world8_stress.json", False), "config": ("config.json", False),
}


def run(name, jsn, brute):
    with tempfile.TemporaryDirectory() as td:
        cpath, ipath = os.path.join(td, "c.bin"), os.path.join(td, "i.bin")
        cmd = [REF_CPU, jsn, "--canvas", cpath, "--ids", ipath] + ([] if brute else ["--ids-bvh-only"])
        out = subprocess.run(cmd, cwd=REF, capture_output=True, text=True, check=True).stdout
        info = json.loads(out.strip().splitlines()[-1])
        raw = np.fromfile(cpath, dtype=np.int32)
        w, h = int(raw[0]), int(raw[1])
        canvas = raw[2:].view(np.uint32).reshape(h, w)
        ids = np.fromfile(ipath, dtype=np.uint8)[8:]
        rec = np.frombuffer(ids.tobytes(), dtype=np.dtype([("trans", "<i4"), ("tri", "<i4"), ("t", "<u4"),
                                                            ("trans_bf", "<i4"), ("tri_bf", "<i4"), ("t_bf", "<u4")]))
        arrays = dict(canvas=canvas, inst=rec["trans"].astype(np.int16), tri=rec["tri"].astype(np.int8),
                      t_bits=rec["t"].copy())
        if brute:
            arrays.update(inst_bf=rec["trans_bf"].astype(np.int16), tri_bf=rec["tri_bf"].astype(np.int8),
                          t_bf_bits=rec["t_bf"].copy())
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrays)
        meta = {k: info[k] for k in ("width", "height", "n_trans", "hash", "nonzero")}
        meta.update(scene=jsn, semantics="serial (rtracer::cpu::update_scene)", ms_frame_1core=info["ms_first"],
                    generator="tests/golden/make_golden.py via oracle/_ref/ref_cpu (unmodified reference sources)")
        json.dump(meta, open(os.path.join(HERE, name + ".json"), "w"), indent=1)
        print(name, meta)


if __name__ == "__main__":
    names = sys.argv[1:] or list(WORLDS)
    for n in names:
        run(n, *WORLDS[n])
