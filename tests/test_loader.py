"""CPU: the C++ host loader (gpu-ray-tracer_b200/host: SceneBuilder / cube_world / Perlin mirror)
produces, for every shipped world, a flattened scene bit-identical to the one the unmodified
reference's loader produces (tests/golden/scene_dumps.json, made by make_scene_dumps.py)."""
import hashlib
import json
import os

import pytest

from conftest import N_INSTANCES, ROOT, scene_path

DUMPS = json.load(open(os.path.join(ROOT, "tests", "golden", "scene_dumps.json")))


@pytest.mark.parametrize("name", sorted(DUMPS))
def test_loader_output_equals_reference_loader(pkg, tmp_path, name):
    w = pkg.World(scene_path(name))
    assert w.n_instances == N_INSTANCES[name] == DUMPS[name]["n_trans"]
    out = tmp_path / "scene.txt"
    w.dump(out)
    data = out.read_bytes()
    assert len(data) == DUMPS[name]["bytes"]
    assert hashlib.sha256(data).hexdigest() == DUMPS[name]["sha256"]


def test_overrides_and_errors(pkg, tmp_path):
    w = pkg.World(scene_path("world4"), width=1920, height=1080, depth=5)
    assert (w.width, w.height, w.desc.env.depth) == (1920, 1080, 5)
    assert w.desc.camera.canvas_width == 1920.0
    g = pkg.World(scene_path("world4"), grid_size=8)
    assert g.n_instances > w.n_instances
    with pytest.raises(RuntimeError):
        pkg.World(str(tmp_path / "missing.json"))
    with pytest.raises(RuntimeError):
        pkg.World(text="{ not json")
    # defaults of the reference's loader (src/procedural/cube_world.cc:44-61): 640x480, grid 8
    d = pkg.World(text='{"atlas": "x", "cubes": [{}], "lights": {}}')
    assert (d.width, d.height) == (640, 480)
