"""CPU: the engine's __host__ __device__ code (BVH8 traversal, conservative triangle filter, exact
candidate test, shading, bounce state machine) compiled for the host (tests/emu) and checked
against the oracle. Catches logic errors without a GPU; the GPU kernels run the same headers."""
import numpy as np
import pytest

import emu_binding as eb
import oracle_binding as ob
from conftest import scene_path

pytestmark = pytest.mark.skipif(not eb.available(), reason="tests/emu/libemu.so not built (run __graft_entry__.build())")


def test_identity_rotation_fast_path_is_bit_exact():
    rng = np.random.default_rng(0)
    v = (rng.standard_normal((200000, 3)) * rng.choice([1e-6, 1e-3, 1, 30, 1e4], (200000, 1))).astype(np.float32)
    v[:10] = 0
    assert eb.lib().emu_check_identity_rotation(v.ctypes.data, len(v)) == 0


@pytest.mark.parametrize("name", ["world1", "world4", "world16", "world8_stress"])
def test_bvh8_traversal_primary_bit_exact(pkg, name):
    w = pkg.World(scene_path(name))
    bvh = ob.OracleBvh(w)
    o = ob.OracleScene(w)
    oi, ot, otime = o.trace_primary(use_bvh=1)
    inst, tri, t, st = eb.trace_primary(w, bvh, 1)
    assert (inst == oi).all() and (tri == ot).all() and (t.view(np.uint32) == otime.view(np.uint32)).all()
    assert st["nodes"] > 0 and st["exact"] <= st["prims"]


# config: Ks = 0 with alpha = 0.8 (phong's specular shortcut: pow * (+-0) is +-0 bit for bit); world8_stress: Ks != 0
@pytest.mark.parametrize("name,sem", [("world1", 0), ("world1", 1), ("world4", 0), ("world8_stress", 0), ("world16", 0),
                                      ("config", 0), ("config", 1)])
def test_render_bit_exact(pkg, name, sem):
    w = pkg.World(scene_path(name), width=320, height=240)
    bvh = ob.OracleBvh(w)
    oc, ost = ob.OracleScene(w).render(semantics=sem, use_bvh=1)
    c, rays = eb.render(w, bvh, 1, sem)
    assert (c == oc).all()
    assert rays["primary"] == ost["rays_primary"] and rays["shadow"] == ost["rays_shadow"]
    assert rays["secondary"] == ost["rays_secondary"]


def test_mixed_materials_gpu_semantics(pkg):
    """Reflective + refractive layers, depth 4: the stale-Isect rules of propagate_ray."""
    w = pkg.World(scene_path("grid64_mixed"), width=160, height=120, grid_size=16, depth=4)
    bvh = ob.OracleBvh(w)
    oc, ost = ob.OracleScene(w).render(semantics=0, use_bvh=1)
    c, rays = eb.render(w, bvh, 1, 0)
    assert (c == oc).all() and rays["secondary"] == ost["rays_secondary"] and rays["secondary"] > 0


def test_edge_and_degenerate_rays(pkg):
    w = pkg.World(scene_path("world8"))
    bvh = ob.OracleBvh(w)
    o = ob.OracleScene(w)
    rng = np.random.default_rng(3)
    n = 6000
    org = rng.uniform(-6, 6, (n, 3)).astype(np.float32)
    org[:, 1] = rng.uniform(0, 14, n)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    boxes = o.instance_boxes()
    k = n // 2
    pick = rng.integers(0, len(boxes), k)
    corner = np.where(rng.integers(0, 2, (k, 3)) == 1, boxes[pick, 3:], boxes[pick, :3])
    d[:k] = corner - org[:k]
    d[k:k + 200] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 200)] * rng.choice([-1, 1], (200, 1))
    d[k + 200:k + 210] = 0
    oi, ot, otime = o.cast_rays(org, d, use_bvh=1)
    gi, gt, gtime = eb.cast_rays(w, bvh, 1, org, d)
    assert (gi == oi).all() and (gt == ot).all() and (gtime.view(np.uint32) == otime.view(np.uint32)).all()
