"""CPU: an independent check of the scene-file reader (gpu-ray-tracer_b200/host/mini_json.h).

The reference builds under oracle/_ref read their scene files through a RapidJSON-API shim that sits on the SAME
mini_json.h as the product loader (caveat), so the golden scene dumps cannot catch a number- or
structure-parsing fault of that reader: both sides would share it. Here every numeric leaf of every shipped world is
parsed a second time by Python's json module (its float() is correctly rounded, as strtod is) and carried through the
loader's float32 arithmetic with numpy (reference: src/procedural/cube_world.cc:39-61,82-138,177-191 — colour terms
x 1/255, Kt / Kr / alpha / eta / ambience / attenuation terms as written, fov in degrees). The result must equal, bit
for bit, what the loader put into the RtSceneDesc."""
import glob
import json
import os

import numpy as np
import pytest

from conftest import ROOT

SCENES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(ROOT, "scenes", "*.json")))
F = np.float32
INV255 = F(1.0) / F(255.0)


def bits(values):
    return np.asarray(values, dtype=np.float32).view(np.uint32).tolist()


def colour(entry, key, scale):
    """A four-component material / light term as the loader stores it: float(value) [* 1/255], None when absent."""
    if key not in entry:
        return None
    v = np.asarray([F(x) for x in entry[key]], dtype=np.float32)
    assert v.shape == (4,)
    return (INV255 * v if scale else v).astype(np.float32)


@pytest.mark.parametrize("name", SCENES)
def test_numeric_leaves_match_python_json(pkg, name):
    path = os.path.join(ROOT, "scenes", name + ".json")
    doc = json.load(open(path))
    w = pkg.World(path)
    d = w.desc

    # top-level scalars
    assert d.width == doc.get("width", 640) and d.height == doc.get("height", 480)
    if "depth" in doc:
        assert d.env.depth == doc["depth"]
    if "ambience" in doc:
        assert bits(list(d.env.ambience)) == bits([F(x) for x in doc["ambience"]])
    if "distance_attenuation" in doc:
        a = doc["distance_attenuation"]
        want = [F(a["constant_term"]), F(a["linear_term"]), F(a["quadratic_term"])]
        assert bits(list(d.env.dist_atten)) == bits(want)
    # camera: unit length and the canvas size go in as written; near = 0.5 W / unit / tan(fov) needs fov only
    assert bits([d.camera.unit_to_pixels]) == bits([F(doc.get("unit_length", 200))])
    assert bits([d.camera.canvas_width, d.camera.canvas_height]) == bits([F(d.width), F(d.height)])

    # one hitable per "cubes" entry, then one per "spheres" entry, in file order; every triangle of a cube carries
    # the entry's material
    entries = list(doc.get("cubes", [])) + list(doc.get("spheres", []))
    assert d.n_meshes == len(entries)
    for m, entry in enumerate(entries):
        mesh = d.meshes[m]
        assert mesh.tri_end > mesh.tri_begin
        for t in (mesh.tri_begin, mesh.tri_end - 1):
            mat = d.triangles[t].mat
            for key, scale in (("Ke", True), ("Ka", True), ("Kd", True), ("Ks", True), ("Kt", False), ("Kr", False)):
                want = colour(entry, key, scale)
                if want is not None:
                    assert bits(list(getattr(mat, key))) == bits(want), (name, m, key)
            for key in ("alpha", "eta"):
                if key in entry:
                    assert bits([getattr(mat, key)]) == bits([F(entry[key])]), (name, m, key)
        if "radius" in entry:
            assert mesh.kind == 1  # RT_HITABLE_SPHERE
            assert bits([mesh.radius]) == bits([F(entry["radius"])])

    # lights: point lights first, then directional ones (src/scene_builder.cu:61-81); a directional light's vector is
    # normalised by the builder, so only its colour is a leaf of the file
    lights = doc.get("lights", {})
    points, dirs = lights.get("point", []), lights.get("directional", [])
    assert d.n_lights == len(points) + len(dirs)
    for i, p in enumerate(points):
        l = d.lights[i]
        assert l.kind == 0
        assert bits(list(l.v)) == bits([F(x) for x in p["pos"]])
        assert bits(list(l.color)) == bits(colour(p, "col", True))
    for i, q in enumerate(dirs):
        l = d.lights[len(points) + i]
        assert l.kind == 1
        assert bits(list(l.color)) == bits(colour(q, "col", True))
        v = np.asarray([float(x) for x in q["dir"]])
        got = np.asarray(list(l.v), dtype=np.float64)
        assert np.allclose(got, v / np.linalg.norm(v), atol=1e-6)

    # sphere instances sit at "pos" after the cubes' transformations
    for k, entry in enumerate(doc.get("spheres", [])):
        inst = d.instances[d.n_instances - len(doc["spheres"]) + k]
        assert bits(list(inst.pose.pos)) == bits([F(x) for x in entry["pos"]])
    w.close()


def test_number_forms(pkg):
    """Literal forms the shipped worlds do not use: exponents, negative zero, long decimals, integers with a sign."""
    text = ('{"atlas": "x", "cubes": [{"Kt": [1e-3, -0.0, 0.30000000000000004, 1.5E+1], "alpha": 2.5e0, "eta": 1.3330001}],\n'
            ' "lights": {"point": [{"pos": [-1e1, 7.25, 123456.789], "col": [255, 0, 127, 255]}]},\t'
            ' "ambience": [0.1, 0.2, 0.7, 1], "depth": 7, "width": 64, "height": 48}\r\n')
    assert json.loads(text)["cubes"][0]["Kt"][3] == 15.0
    w = pkg.World(text=text)
    d = w.desc
    mat = d.triangles[d.meshes[0].tri_begin].mat
    assert bits(list(mat.Kt)) == bits([F(1e-3), F(-0.0), F(0.30000000000000004), F(15.0)])
    assert bits([mat.alpha, mat.eta]) == bits([F(2.5), F(1.3330001)])
    assert bits(list(d.lights[0].v)) == bits([F(-10.0), F(7.25), F(123456.789)])
    assert bits(list(d.lights[0].color)) == bits(INV255 * np.asarray([255, 0, 127, 255], dtype=np.float32))
    assert bits(list(d.env.ambience)) == bits([F(0.1), F(0.2), F(0.7), F(1)])
    assert (d.env.depth, d.width, d.height) == (7, 64, 48)
    w.close()
