"""CPU: the warp-coherent traversal of the default pipeline (csrc/rt_beam.cuh: beam_query, the walk of k_tile<1> and
k_shadow_beam) executed on emulated warps — 32 fibers per warp, collectives as rendezvous (tests/emu/emu_warp.cu) — and
checked against the oracle. The header is the one the kernels compile: shuffles, ballots, the warp-wide OR, the stack spread
over the lanes, the octant fallback and the deferred exact tests all run as written. Catches logic errors without a GPU."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import emu_binding as eb
import oracle_binding as ob
from conftest import ROOT, scene_path

pytestmark = pytest.mark.skipif(not eb.available(), reason="tests/emu/libemu.so not built (run __graft_entry__.build())")


def bits(a):
    return np.asarray(a, np.float32).view(np.uint32)


def test_warp_stand_ins():
    """prmt selector semantics, ffs / clz / popc, and shuffles / ballots / reductions among fibers that drift apart."""
    assert eb.lib().emu_warp_selftest() == 0


@pytest.mark.parametrize("name,kw", [("world1", {}), ("world4", dict(width=321, height=243)), ("world16", dict(width=320, height=240)),
                                     ("world8_stress", dict(width=320, height=240)), ("spheres8", dict(width=320, height=240)),
                                     ("textured8", dict(width=160, height=120)), ("grid64_mirror", dict(width=160, height=120, grid_size=24))])
def test_beam_primary_bit_exact(pkg, name, kw):
    """Camera rays, one emulated warp per 8x4 tile (ragged canvases leave inactive lanes): ids, triangle and t bits of the
    oracle; a node is fetched once per warp, so far fewer fetches than per-ray node visits."""
    w = pkg.World(scene_path(name), **kw)
    bvh = ob.OracleBvh(w)
    oi, ot, otime = ob.OracleScene(w).trace_primary(use_bvh=1)
    inst, tri, t, st = eb.beam_primary(w, bvh)
    assert (inst == oi).all() and (tri == ot).all() and (bits(t) == bits(otime)).all()
    li, lt, ltime, lst = eb.trace_primary(w, bvh, 1)  # the per-lane walk (bvh_query) over the same tree
    assert (inst == li).all() and (bits(t) == bits(ltime)).all()
    assert 0 < st["node_fetches"] * 8 < st["nodes"] and st["exact"] <= st["prims"]
    # without the leaf-box rule (BVH8 only)
    oi2, ot2, otime2 = ob.OracleScene(w).trace_primary(use_bvh=2) if name == "world4" else (None, None, None)
    if oi2 is not None:
        i2, t2, tt2, _ = eb.beam_primary(w, bvh, use_box=0)
        assert (i2 == oi2).all() and (t2 == ot2).all() and (bits(tt2) == bits(otime2)).all()


@pytest.mark.parametrize("name", ["world4", "world8", "world8_stress"])
def test_beam_primary_on_moved_scenes(pkg, name):
    """Camera moved and cubes rotated in place (tests/golden/moved_*.json, made by the unmodified reference): the general
    quaternion path of the exact test behind the warp-coherent walk; equals the reference's own ids and t."""
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "moved_%s.json" % name)))
    g = np.load(os.path.join(ROOT, "tests", "golden", "moved_%s.npz" % name))
    w = pkg.World(scene_path(name))
    w.apply_moves(meta["moves"])
    bvh = ob.OracleBvh(w)
    inst, tri, t, _ = eb.beam_primary(w, bvh)
    assert (inst == g["inst"]).all() and (tri == g["tri"]).all() and (bits(t) == g["t_bits"]).all()


@pytest.mark.parametrize("name,kw", [("world4", dict(width=321, height=243)), ("world16", dict(width=320, height=240)),
                                     ("world8_stress", dict(width=320, height=240)), ("spheres8", dict(width=160, height=120))])
@pytest.mark.parametrize("cull", [0, 1])
def test_beam_shadow_equals_per_lane_walk(pkg, name, kw, cull):
    """The occlusion query of k_shadow_beam / k_tile (one walk per tile and light, early exit when every ray of the warp is
    answered) gives every shadow ray the flag of the per-lane walk of k_shadow — with and without the zero-weight culling."""
    w = pkg.World(scene_path(name), **kw)
    bvh = ob.OracleBvh(w)
    beam, lane = eb.beam_shadow(w, bvh, cull)
    assert (beam == lane).all()
    traced = beam != 255
    assert traced.sum() > 500 and (beam[traced] == 1).any() and (beam[traced] == 0).any()
    if cull:
        full, _ = eb.beam_shadow(w, bvh, 0)
        assert traced.sum() <= (full != 255).sum() and (beam[traced] == full[traced]).all()


def test_beam_cast_rays_mixed_octants_and_degenerate(pkg):
    """Rays that are not a screen tile: random origins and directions. Unsorted, the 32 rays of a warp span several
    direction octants (the per-lane fallback); sorted by octant, every warp shares one and walks as a beam whose origin box
    is the whole world. Corners of instance boxes, axis-parallel and zero directions included. Oracle ids and t bits."""
    w = pkg.World(scene_path("world8"))
    bvh = ob.OracleBvh(w)
    o = ob.OracleScene(w)
    rng = np.random.default_rng(11)
    n = 4096 + 17
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
    for order in (np.arange(n), np.argsort((d[:, 0] < 0) * 1 + (d[:, 1] < 0) * 2 + (d[:, 2] < 0) * 4, kind="stable")):
        oo, dd = org[order], d[order]
        oi, ot, otime = o.cast_rays(oo, dd, use_bvh=1)
        gi, gt, gtime = eb.beam_cast_rays(w, bvh, oo, dd)
        assert (gi == oi).all() and (gt == ot).all() and (bits(gtime) == bits(otime)).all()
        assert (gi >= 0).sum() > 500


def test_beam_on_a_deep_tree_of_coincident_triangles(pkg):
    """512 coincident instances = 6 144 triangles over 12 Morton codes: leaves full of duplicates, a BVH8 several levels deep;
    the warp's register stack and the candidate rounds hold. Hit / miss, triangle and t of the oracle (instance ids of exact
    ties go to the lowest index)."""
    n = 512
    w = pkg.World(scene_path("world1"), width=160, height=120, depth=0)
    inst = (pkg.RtInstance * n)()
    for i in range(n):
        C.memmove(C.byref(inst[i]), C.byref(w.desc.instances[0]), C.sizeof(pkg.RtInstance))
    w.desc.instances = C.cast(inst, C.POINTER(pkg.RtInstance))
    w.desc.n_instances = n
    bvh = ob.OracleBvh(w)
    assert bvh.n_prims == 12 * n and bvh.levels >= 3
    oi, ot, otime = ob.OracleScene(w).trace_primary(use_bvh=1)
    gi, gt, gtime, _ = eb.beam_primary(w, bvh)
    li, lt, ltime, _ = eb.trace_primary(w, bvh, 1)
    assert ((gi >= 0) == (oi >= 0)).all() and (gt == ot).all() and (bits(gtime) == bits(otime)).all()
    assert (gi == li).all() and (gi >= 0).sum() > 50


@pytest.mark.parametrize("name", ["world4", "world16", "world8_stress"])
def test_beam_primary_tolerance_mode(pkg, name):
    """The opt-in tolerance mode of the same walk (beam_query<.., FAST>: hits clearly inside a triangle skip the
    reference-exact test): north_star's bars against the oracle — ids on >= 99.99 % of the rays, t within 1e-4 relative —
    and it really skips exact tests."""
    w = pkg.World(scene_path(name), width=320, height=240)
    bvh = ob.OracleBvh(w)
    oi, ot, otime = ob.OracleScene(w).trace_primary(use_bvh=1)
    inst, tri, t, st = eb.beam_primary(w, bvh, use_box=2)
    _, _, _, st_exact = eb.beam_primary(w, bvh, use_box=1)
    same = (inst == oi) & (tri == ot)
    hit = same & (oi >= 0)
    rel = np.abs(t[hit].astype(np.float64) - otime[hit]) / np.abs(otime[hit])
    assert same.mean() >= 0.9999 and rel.max() <= 1e-4
    assert st["exact"] < 0.7 * st_exact["exact"]
