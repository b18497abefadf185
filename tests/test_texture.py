"""Texture-atlas sampling. The reference leaves it as `// TODO: implement texture mapping`
(src/rayprimitives/phong.cu:19-23), so there is nothing of the reference to pin against: the oracle's restatement of
the intended lookup (TextureCoords + Isect::uv -> point-sampled, clamped texel, oracle/rt_oracle.c diffuse_colour) is
checked through properties here, and the engine against the oracle bit for bit."""
import json
import os

import numpy as np
import pytest

import emu_binding as eb
import oracle_binding as ob
from conftest import scene_path

needs_emu = pytest.mark.skipif(not eb.available(), reason="tests/emu/libemu.so not built (run __graft_entry__.build())")

KD0, KD1 = [210, 105, 0, 255], [124, 252, 0, 255]


def write_ppm(path, img):
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(img.astype(np.uint8).tobytes())


def two_tile_world(tmp_path, textured, atlas=None, tiles=((0, 0, 16), (16, 0, 16))):
    cfg = json.load(open(scene_path("world8")))
    cfg["cubes"][0]["Kd"], cfg["cubes"][1]["Kd"] = KD0, KD1
    cfg["width"], cfg["height"] = 240, 180
    if textured:
        cfg["cubes"][0]["texture"], cfg["cubes"][1]["texture"] = list(tiles[0]), list(tiles[1])
        p = os.path.join(str(tmp_path), "atlas.ppm")
        write_ppm(p, atlas)
        cfg["atlas_ppm"] = p
    return cfg


def test_uniform_tiles_reproduce_the_untextured_frame(pkg, tmp_path):
    """An atlas whose two tiles are filled with the cubes' own Kd gives the untextured frame bit for bit: the lookup
    replaces Kd in the diffuse term and nothing else (alpha of a PPM texel is 1 = Kd's 255 / 255)."""
    atlas = np.zeros((16, 32, 3), np.uint8)
    atlas[:, :16] = KD0[:3]
    atlas[:, 16:] = KD1[:3]
    plain = pkg.World(text=json.dumps(two_tile_world(tmp_path, False)))
    tex = pkg.World(text=json.dumps(two_tile_world(tmp_path, True, atlas)))
    assert tex.desc.atlas_width == 32 and tex.desc.atlas_height == 16 and tex.desc.triangles[0].tex_degenerate == 0
    a, sa = ob.OracleScene(plain).render(semantics=0, use_bvh=1)
    b, sb = ob.OracleScene(tex).render(semantics=0, use_bvh=1)
    assert (a == b).all() and sa["rays_total"] == sb["rays_total"]
    # a different texel colour does change the frame, and only where the diffuse term is lit
    atlas[:, 16:] = [10, 20, 200]
    c, _ = ob.OracleScene(pkg.World(text=json.dumps(two_tile_world(tmp_path, True, atlas)))).render(semantics=0, use_bvh=1)
    assert (c != a).any() and ((c != a) <= (a != 0)).all()


def test_texture_coordinates_follow_the_cube_faces(pkg):
    """build_textured_cube: every triangle's three vertices land on corners of its tile, the two triangles of a face
    cover the tile once (|u x v| = tile area), origins are integers inside the atlas."""
    w = pkg.World(scene_path("textured8"))
    d = w.desc
    assert d.atlas_width == 64 and d.atlas_height == 64
    for t in range(24):  # two textured cube meshes
        tr = d.triangles[t]
        ox, oy, S = (0, 0, 16) if t < 12 else (16, 32, 16)
        p0 = np.array([tr.tex_x, tr.tex_y], float)
        p1, p2 = p0 + np.array(tr.tex_u[:]), p0 + np.array(tr.tex_v[:])
        for p in (p0, p1, p2):
            assert p[0] in (ox, ox + S) and p[1] in (oy, oy + S)
        assert abs(tr.tex_u[0] * tr.tex_v[1] - tr.tex_u[1] * tr.tex_v[0]) == S * S
    assert d.triangles[24].tex_degenerate == 1  # the sphere's material record


@needs_emu
@pytest.mark.parametrize("tiles", [((0, 0, 16), (16, 32, 16)), ((56, 56, 16), (-8, -8, 16))], ids=["inside", "clamped"])
def test_engine_code_bit_exact_on_textures(pkg, tmp_path, tiles):
    """Engine shading code (host build of the same headers) == oracle, also when tiles hang over the atlas edge and
    the lookup clamps."""
    cfg = json.load(open(scene_path("textured8")))
    cfg["atlas_ppm"] = scene_path("atlas64").replace(".json", ".ppm")
    cfg["cubes"][0]["texture"], cfg["cubes"][1]["texture"] = list(tiles[0]), list(tiles[1])
    cfg["width"], cfg["height"] = 240, 180
    w = pkg.World(text=json.dumps(cfg))
    oc, ost = ob.OracleScene(w).render(semantics=0, use_bvh=1)
    c, rays = eb.render(w, ob.OracleBvh(w), 1, 0)
    assert (c == oc).all(), int((c != oc).sum())
    assert len(np.unique(oc)) > 100  # the checker tiles are in the frame


@pytest.mark.gpu
def test_gpu_textures_bit_exact(pkg, gpu_ok):
    w = pkg.World(scene_path("textured8"))
    s = pkg.Scene(w)
    o = ob.OracleScene(w)
    oc, ost = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    for flags in (pkg.RT_RENDER_REBUILD_BVH, pkg.RT_RENDER_FUSED_SHADE, pkg.RT_RENDER_TRACE_UNLIT):
        c, st = s.render(flags=flags)
        assert (c == oc).all(), (flags, int((c != oc).sum()))
        assert st["rays_total"] == ost["rays_total"]
    c2, _ = s.render(flags=pkg.RT_RENDER_NO_BVH)
    oc2, _ = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=0)
    assert (c2 == oc2).all()
    c3, _ = s.render(semantics=pkg.RT_SEMANTICS_CPU)  # no Kr / Kt in this scene: the serial semantics terminate
    oc3, _ = o.render(semantics=pkg.RT_SEMANTICS_CPU, use_bvh=1)
    assert (c3 == oc3).all()
    s.close()


@pytest.mark.gpu
def test_gpu_textured_mirror_bounces(pkg, gpu_ok, tmp_path):
    """Textured cubes seen through bounces: the mirror layer of world8_stress reflects textured layers."""
    cfg = json.load(open(scene_path("world8_stress")))
    cfg["cubes"][0]["texture"], cfg["cubes"][1]["texture"] = [0, 16, 16], [32, 48, 16]
    cfg["atlas_ppm"] = scene_path("atlas64").replace(".json", ".ppm")
    cfg["width"], cfg["height"], cfg["depth"] = 320, 240, 4
    w = pkg.World(text=json.dumps(cfg))
    s = pkg.Scene(w)
    oc, ost = ob.OracleScene(w).render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    c, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH)
    assert (c == oc).all() and st["rays_secondary"] == ost["rays_secondary"] > 0
    c, st = s.render(flags=pkg.RT_RENDER_FUSED_SHADE)
    assert (c == oc).all()
    s.close()
