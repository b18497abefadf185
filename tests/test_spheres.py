"""Analytic sphere hitables. The reference has no second Hitable subclass, so there is no
reference output to pin: the oracle's Sphere::hit_local (oracle/rt_oracle.c, written in the reference's
conventions) is checked against closed-form float64 geometry here, and the engine against the oracle bit for bit
(CPU: the engine's __host__ __device__ code through tests/emu; -m gpu: the CUDA kernels through the C ABI)."""
import json

import numpy as np
import pytest

import emu_binding as eb
import oracle_binding as ob
from conftest import scene_path

needs_emu = pytest.mark.skipif(not eb.available(), reason="tests/emu/libemu.so not built (run __graft_entry__.build())")


def spheres_of(name="spheres8"):
    cfg = json.load(open(scene_path(name)))
    return [(np.array(s["pos"], np.float64), float(s["radius"])) for s in cfg["spheres"]]


def analytic_nearest(org, d, spheres):
    """float64 closest positive root over all spheres; returns (index or -1, t)."""
    best_t = np.full(len(org), np.inf)
    best_i = np.full(len(org), -1)
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    for i, (c, r) in enumerate(spheres):
        oc = org - c
        b = (oc * d).sum(1)
        cc = (oc * oc).sum(1) - r * r
        disc = b * b - cc
        ok = disc >= 0
        sq = np.sqrt(np.where(ok, disc, 0))
        t0, t1 = -b - sq, -b + sq
        t = np.where(t0 >= 1e-5, t0, t1)
        ok &= (t >= 1e-5) & (t < best_t)
        best_t = np.where(ok, t, best_t)
        best_i = np.where(ok, i, best_i)
    return best_i, best_t


def test_oracle_sphere_hits_match_closed_form(pkg):
    """Rays against the five spheres alone (cubes removed): instance ids equal the float64 solution and t agrees to 1e-4
    relative (north_star's tolerance for hit distances), from outside, from inside and for misses."""
    cfg = json.load(open(scene_path("spheres8")))
    cfg["cubes"] = []
    w = pkg.World(text=json.dumps(cfg))
    assert w.n_instances == 5 and w.desc.n_meshes == 5
    o = ob.OracleScene(w)
    sph = spheres_of()
    rng = np.random.default_rng(7)
    n = 20000
    org = rng.uniform(-8, 8, (n, 3)); org[:, 1] = rng.uniform(2, 14, n)
    k = n // 4  # a quarter of the origins inside a sphere
    pick = rng.integers(0, len(sph), k)
    u = rng.standard_normal((k, 3)); u /= np.linalg.norm(u, axis=1, keepdims=True)
    org[:k] = np.array([sph[i][0] for i in pick]) + u * (np.array([sph[i][1] for i in pick]) * rng.uniform(0, 0.9, k))[:, None]
    tgt = np.array([sph[i][0] for i in rng.integers(0, len(sph), n)]) + rng.standard_normal((n, 3)) * 1.2
    d = tgt - org
    org32, d32 = org.astype(np.float32), d.astype(np.float32)
    inst, tri, t = o.cast_rays(org32, d32, use_bvh=1)
    ai, at = analytic_nearest(org32.astype(np.float64), d32.astype(np.float64), sph)
    # grazing rays may flip between hit and miss in fp32: compare where the float64 discriminant is not marginal
    sure = np.ones(n, bool)
    dn = d32.astype(np.float64); dn /= np.linalg.norm(dn, axis=1, keepdims=True)
    for c, r in sph:
        oc = org32.astype(np.float64) - c
        b = (oc * dn).sum(1)
        disc = b * b - ((oc * oc).sum(1) - r * r)
        sure &= np.abs(disc) > 1e-3
    assert sure.mean() > 0.95
    assert (inst[sure] == ai[sure]).all()
    hit = sure & (ai >= 0)
    assert hit.sum() > n // 4 and (tri[hit] == 0).all()
    assert np.allclose(t[hit], at[hit], rtol=1e-4, atol=1e-5)
    brute = o.cast_rays(org32, d32, use_bvh=0)
    assert (brute[0] == inst).all() and (brute[2].view(np.uint32) == t.view(np.uint32)).all()


def test_sphere_normals_are_radial(pkg):
    w = pkg.World(scene_path("spheres8"), width=96, height=72)
    o = ob.OracleScene(w)
    inst, tri, t = o.trace_primary(use_bvh=1)
    sph = spheres_of()
    first = w.n_instances - len(sph)
    ys, xs = np.divmod(np.nonzero(inst >= first)[0], 96)
    assert len(xs) > 200
    for x, y in list(zip(xs, ys))[::7]:
        hit = pkg.RtHit()
        ob.lib().or_debug_cast(o._h, 0, 1, int(x), int(y), pkg.C.byref(hit))
        org, d = o.primary_ray(int(x), int(y))
        c, r = sph[hit.instance - first]
        p = org.astype(np.float64) + hit.t * d.astype(np.float64)
        assert np.allclose(np.array(hit.normal[:]), (p - c) / r, atol=2e-4)


def test_loader_spheres_and_validation(pkg):
    w = pkg.World(scene_path("spheres8"))
    d = w.desc
    kinds = [d.meshes[i].kind for i in range(d.n_meshes)]
    assert kinds == [0, 0] + [pkg.RT_HITABLE_SPHERE] * 5
    for i in range(2, 7):
        assert d.meshes[i].tri_end - d.meshes[i].tri_begin == 1 and d.meshes[i].radius > 0
    assert w.n_instances == 380 + 5  # world8's cubes + the spheres


@needs_emu
def test_engine_code_bit_exact_on_spheres(pkg):
    """BVH8 built over cubes + spheres by the host build, walked by the engine's traversal; ids, t and the rendered
    frame (matte, mirror, glass and emissive spheres, depth 5) equal the oracle's bit for bit."""
    w = pkg.World(scene_path("spheres8"), width=240, height=180)
    bvh = ob.OracleBvh(w)
    assert bvh.n_prims == 380 * 12 + 5
    o = ob.OracleScene(w)
    oi, ot, otime = o.trace_primary(use_bvh=1)
    inst, tri, t, st = eb.trace_primary(w, bvh, 1)
    assert (inst == oi).all() and (tri == ot).all() and (t.view(np.uint32) == otime.view(np.uint32)).all()
    assert (oi >= 380).sum() > 1000
    # CUDA semantics only: the reference's serial propagate_helper never decrements towards its depth guard
    # (SURVEY.md §3.4), so a ray refracted into a closed glass sphere recurses until the stack is gone
    oc, ost = o.render(semantics=0, use_bvh=1)
    c, rays = eb.render(w, bvh, 1, 0)
    assert (c == oc).all(), int((c != oc).sum())
    assert rays["secondary"] == ost["rays_secondary"] and rays["secondary"] > 0 and rays["shadow"] == ost["rays_shadow"]


@needs_emu
def test_engine_code_random_rays_on_spheres(pkg):
    w = pkg.World(scene_path("spheres8"))
    bvh = ob.OracleBvh(w)
    o = ob.OracleScene(w)
    rng = np.random.default_rng(11)
    n = 8000
    org = rng.uniform(-6, 6, (n, 3)).astype(np.float32); org[:, 1] = rng.uniform(2, 14, n).astype(np.float32)
    sph = spheres_of()
    tgt = np.array([sph[i][0] for i in rng.integers(0, len(sph), n)]) + rng.standard_normal((n, 3)) * 1.5
    d = (tgt - org).astype(np.float32)
    org[:50] = np.array(sph[1][0], np.float32)  # from the centre of the mirror sphere
    oi, ot, otime = o.cast_rays(org, d, use_bvh=1)
    gi, gt, gtime = eb.cast_rays(w, bvh, 1, org, d)
    assert (gi == oi).all() and (gt == ot).all() and (gtime.view(np.uint32) == otime.view(np.uint32)).all()
    assert (oi >= 380).mean() > 0.2


@pytest.mark.gpu
def test_gpu_spheres_bit_exact(pkg, gpu_ok):
    w = pkg.World(scene_path("spheres8"))
    s = pkg.Scene(w)
    o = ob.OracleScene(w)
    # build
    s.build_bvh()
    d = s.download_bvh()
    hb = ob.OracleBvh(w)
    assert d["n_prims"] == hb.n_prims and d["n_nodes8"] == hb.n_nodes8
    assert (d["morton"] == hb.morton).all() and (d["perm"] == hb.perm).all()
    assert (d["node_boxes"].view(np.uint32) == hb.node_boxes.view(np.uint32)).all()
    assert (d["nodes8"] == hb.nodes8).all() and (d["prims8"].view(np.uint32) == hb.prims8.view(np.uint32)).all()
    # closest hits
    oi, ot, otime = o.trace_primary(use_bvh=1)
    inst, tri, t = s.trace_primary(use_bvh=1)
    assert (inst == oi).all() and (tri == ot).all() and (t.view(np.uint32) == otime.view(np.uint32)).all()
    # frames: wavefront (split and fused shading), serial semantics, brute force
    oc, ost = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    for flags in (pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_STATS, pkg.RT_RENDER_FUSED_SHADE, pkg.RT_RENDER_REFIT_BVH):
        c, st = s.render(flags=flags)
        assert (c == oc).all(), (flags, int((c != oc).sum()))
        assert st["rays_total"] == ost["rays_total"] and st["rays_secondary"] == ost["rays_secondary"] > 0
    c1, _ = s.render(flags=pkg.RT_RENDER_NO_BVH)  # one thread per pixel over brute force (`-r`)
    oc1, _ = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=0)
    assert (c1 == oc1).all()
    hit = s.debug_cast(300, 200)
    ohit = pkg.RtHit()
    ob.lib().or_debug_cast(o._h, 0, 1, 300, 200, pkg.C.byref(ohit))
    assert hit.instance == ohit.instance and list(hit.normal) == list(ohit.normal) and list(hit.color) == list(ohit.color)
    s.close()


@pytest.mark.gpu
def test_gpu_spheres_only_scene_and_moved_spheres(pkg, gpu_ok):
    """No triangle mesh at all (n_vertices == 0), then spheres moved and the tree refitted."""
    cfg = json.load(open(scene_path("spheres8")))
    cfg["cubes"] = []
    w = pkg.World(text=json.dumps(cfg), width=320, height=240)
    assert w.desc.n_vertices == 0
    s = pkg.Scene(w)
    o = ob.OracleScene(w)
    c, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH)
    oc, ost = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    assert (c == oc).all() and st["rays_total"] == ost["rays_total"]
    w.apply_moves("g:0:0.5:0.25:-0.5,g:1:-1:0:0.5,l:3:0:1.5:0,r:2:0:1:0:0.7")
    s.set_instances(w.desc.instances, w.desc.n_instances)
    o2 = ob.OracleScene(w)
    oc2, _ = o2.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    c2, _ = s.render(flags=pkg.RT_RENDER_REFIT_BVH)
    c3, _ = s.render(flags=pkg.RT_RENDER_REBUILD_BVH)
    assert (oc2 != oc).any() and (c2 == oc2).all() and (c3 == oc2).all()
    s.close()


@pytest.mark.gpu
def test_gpu_sphere_validation(pkg, gpu_ok):
    w = pkg.World(scene_path("spheres8"))
    w.desc.meshes[3].radius = -1.0
    with pytest.raises(RuntimeError):
        pkg.Scene(w)
    w.desc.meshes[3].radius = 1.0
    w.desc.meshes[3].kind = 7
    with pytest.raises(RuntimeError):
        pkg.Scene(w)
