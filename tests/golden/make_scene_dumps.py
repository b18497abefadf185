#!/usr/bin/env python3
"""Writes tests/golden/scene_dumps.json: sha256 of the flattened scene the UNMODIFIED reference's
loader (procedural::cpu::generate -> SceneBuilder) produces for each shipped world, dumped bit for
bit (hex floats) by oracle/_ref/ref_cpu --scene. Only runs in the build container."""
import hashlib
import json
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
REF_CPU = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")
WORLDS = ["world1", "world2", "world4", "world8", "world16", "world8_stress", "config"]

out = {}
for name in WORLDS:
    with tempfile.TemporaryDirectory() as td:
        dump = os.path.join(td, "scene.txt")
        subprocess.run([REF_CPU, name + ".json", "--scene", dump], cwd=REF, check=True, capture_output=True)
        data = open(dump, "rb").read()
        out[name] = {"sha256": hashlib.sha256(data).hexdigest(), "bytes": len(data),
                     "n_trans": int([l for l in data.decode().splitlines() if l.startswith("n_trans")][0].split()[1])}
        print(name, out[name])
json.dump(out, open(os.path.join(HERE, "scene_dumps.json"), "w"), indent=1)
