"""-m gpu: the CUDA engine, called through the C ABI, against the oracle and the golden fixtures
the unmodified reference produced (tests/golden/). Integer ids bit-exact, t bit-exact (tolerance
stated where looser), canvases bit-exact against the oracle."""
import os

import numpy as np
import pytest

import oracle_binding as ob
from conftest import WORLDS, N_INSTANCES, golden, scene_path

pytestmark = pytest.mark.gpu

# north_star tolerances
ID_MATCH_MIN = 0.9999
T_REL_TOL = 1e-4
RMSE_MAX = 1e-3


def rmse(a, b):
    def ch(c):
        return np.stack([(c >> s) & 0xff for s in (24, 16, 8, 0)], -1).astype(np.float64) / 255.0
    return float(np.sqrt(np.mean((ch(a) - ch(b)) ** 2)))


@pytest.fixture(scope="module")
def scenes(pkg, gpu_ok):
    cache = {}

    def get(name, **kw):
        key = (name, tuple(sorted(kw.items())))
        if key not in cache:
            w = pkg.World(scene_path(name), **kw)
            cache[key] = (w, pkg.Scene(w))
        return cache[key]
    yield get
    for w, s in cache.values():
        s.close()


@pytest.mark.parametrize("name", WORLDS)
def test_primary_ids_vs_reference_golden(pkg, scenes, name):
    """Closest-hit (instance, triangle, t) of every primary ray equals what the unmodified
    reference's serial path produced (tests/golden/*.npz)."""
    w, s = scenes(name)
    assert w.n_instances == N_INSTANCES[name]
    meta, g = golden(name)
    inst, tri, t = s.trace_primary(use_bvh=1)
    same = (inst == g["inst"]) & (tri == g["tri"])
    assert same.mean() >= ID_MATCH_MIN
    assert same.all(), f"{(~same).sum()} primary rays differ"
    tg = g["t_bits"].view(np.float32)
    hit = inst >= 0
    assert np.all(np.abs(t[hit] - tg[hit]) <= T_REL_TOL * np.abs(tg[hit]))
    assert (t.view(np.uint32) == g["t_bits"]).all()
    if "inst_bf" in g:
        inst0, tri0, t0 = s.trace_primary(use_bvh=0)  # `-r`: brute force, no leaf-box rule
        assert (inst0 == g["inst_bf"]).all() and (tri0 == g["tri_bf"]).all()
        assert (t0.view(np.uint32) == g["t_bf_bits"]).all()


@pytest.mark.parametrize("name", WORLDS)
def test_tracer_variants_agree(pkg, scenes, name):
    """BVH8 traversal == brute force, with and without the reference's leaf-box rule."""
    w, s = scenes(name)
    a = s.trace_primary(use_bvh=1)
    b = s.trace_primary(use_bvh=3)
    for x, y in zip(a, b):
        assert (x.view(np.uint32) == y.view(np.uint32)).all()
    if w.n_instances <= 400:
        c = s.trace_primary(use_bvh=2)
        d = s.trace_primary(use_bvh=0)
        for x, y in zip(c, d):
            assert (x.view(np.uint32) == y.view(np.uint32)).all()


@pytest.mark.parametrize("name", WORLDS)
def test_canvas_cpu_semantics_vs_reference_golden(pkg, scenes, name):
    """RT_SEMANTICS_CPU frame == the canvas the reference's serial path rendered."""
    w, s = scenes(name)
    meta, g = golden(name)
    canvas, st = s.render(semantics=pkg.RT_SEMANTICS_CPU)
    h, nz = ob.canvas_hash(canvas)
    assert rmse(canvas, g["canvas"]) <= RMSE_MAX
    assert (canvas == g["canvas"]).all()
    assert "%016x" % h == meta["hash"] and nz == meta["nonzero"]


MIXED = dict(width=160, height=120, grid_size=16, depth=4)  # sized for the oracle (seconds)


@pytest.mark.parametrize("name", WORLDS + ["grid64_mixed"])
def test_canvas_gpu_semantics_vs_oracle(pkg, scenes, name):
    """Wavefront pipeline (reference CUDA semantics) == oracle's restatement of propagate_ray."""
    w, s = scenes(name, **(MIXED if name == "grid64_mixed" else {}))
    o = ob.OracleScene(w)
    oc, ost = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    canvas, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_STATS)
    assert rmse(canvas, oc) <= RMSE_MAX
    assert (canvas == oc).all(), f"{(canvas != oc).sum()} pixels differ"
    assert st["rays_primary"] == ost["rays_primary"]
    assert st["rays_shadow"] == ost["rays_shadow"]
    assert st["rays_secondary"] == ost["rays_secondary"]
    assert st["rays_total"] == ost["rays_total"]
    # the one-thread-per-pixel kernel over brute force (`-r`) gives the same frame
    if w.n_instances <= 100:
        c2, _ = s.render(flags=pkg.RT_RENDER_NO_BVH)
        oc2, _ = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=0)
        assert (c2 == oc2).all()
    # update_scene: the reference-shaped call
    c3 = s.update_scene(16, True)
    assert (c3 == oc).all()
    # all-opaque scenes shade through gen / shadow / resolve by default; the fused kernel agrees
    c4, st4 = s.render(flags=pkg.RT_RENDER_FUSED_SHADE)
    assert (c4 == oc).all() and st4["rays_total"] == ost["rays_total"] and st4["rays_culled"] == 0
    # shadow rays of lights with zero Phong weight are answered without a traversal by default;
    # tracing them like the reference does gives the same frame and the same query count
    c5, st5 = s.render(flags=pkg.RT_RENDER_TRACE_UNLIT)
    assert (c5 == oc).all() and st5["rays_total"] == ost["rays_total"] and st5["rays_culled"] == 0
    assert st["rays_culled"] <= st["rays_shadow"]


# (scene, loader overrides): the small worlds sort 32-bit Morton prefixes in 4 passes with the scan fused into the
# scatter; grid64 (256 k triangles) takes the separate per-digit scan; world16 at grid 150 (1.3 M triangles, more than
# 2^20) the 63-bit keys in 8 passes
BVH_CASES = [("world1", {}), ("world4", {}), ("world16", {}), ("world8_stress", {}), ("grid64", {}),
             ("world16", {"grid_size": 150})]


# builder: auto = the one-launch cluster build when the scene fits a cluster (n <= 16 x 4096 triangles), else the
# multi-kernel pipeline; multi = the pipeline forced; cluster8 = an 8-CTA cluster forced
BVH_BUILDS = [(n, kw, b) for n, kw in BVH_CASES for b in (("auto", "multi", "cluster8") if not kw and n != "grid64" else ("auto",))]


@pytest.mark.parametrize("name,kw,builder", BVH_BUILDS, ids=lambda v: v if isinstance(v, str) else "-".join("%s%s" % i for i in v.items()))
def test_bvh_build_bit_exact_vs_host(pkg, scenes, monkeypatch, name, kw, builder):
    """GPU LBVH (Morton, radix sort, Karras, refit) and BVH8 collapse == the sequential host build, for both builders."""
    w, s = scenes(name, **kw)
    monkeypatch.delenv("RT_BUILD_PATH", raising=False)
    monkeypatch.delenv("RT_BUILD_CLUSTER", raising=False)
    if builder == "multi":
        monkeypatch.setenv("RT_BUILD_PATH", "1")
    elif builder == "cluster8":
        monkeypatch.setenv("RT_BUILD_CLUSTER", "8")
    bs = s.build_bvh()
    path, csize = s.debug_get(pkg.RT_DEBUG_BUILD_PATH), s.debug_get(pkg.RT_DEBUG_BUILD_CLUSTER)
    if builder == "multi" or bs["n_prims"] > 16 * 4096:
        assert path == 1 and bs["kernel_launches"] >= 10
    elif builder == "cluster8":
        assert (path, csize, bs["kernel_launches"]) == (2, 8, 1) if bs["n_prims"] <= 8 * 4096 else path == 1
    else:
        assert path == 2 and csize in (8, 16) and bs["kernel_launches"] == 1
    d = s.download_bvh()
    hb = ob.OracleBvh(w)
    assert bs["n_levels8"] == hb.levels
    assert d["n_prims"] == hb.n_prims and d["n_nodes8"] == hb.n_nodes8
    assert (d["bounds"] == hb.bounds).all()
    assert (d["morton"] == hb.morton).all()
    assert (d["perm"] == hb.perm).all()
    assert (d["left"] == hb.left).all() and (d["right"] == hb.right).all()
    assert (d["parent"][1:] == hb.parent[1:]).all()
    assert (d["node_boxes"].view(np.uint32) == hb.node_boxes.view(np.uint32)).all()
    assert (d["prim_order8"] == hb.prim_order8).all()
    assert (d["nodes8"] == hb.nodes8).all()
    assert (d["prims8"].view(np.uint32) == hb.prims8.view(np.uint32)).all()
    # the tree the builder left is the one the frame is traced through, and a refit of it reproduces it
    if name in ("world4", "world16"):
        c, _ = s.render()
        c2, _ = s.render(flags=pkg.RT_RENDER_REFIT_BVH)
        assert (c == c2).all()
        d2 = s.download_bvh()
        assert (d2["nodes8"] == hb.nodes8).all() and (d2["node_boxes"].view(np.uint32) == hb.node_boxes.view(np.uint32)).all()


def test_random_rays_vs_oracle(pkg, scenes):
    """Incoherent rays, rays aimed at triangle edges / vertices, axis-parallel and degenerate rays."""
    w, s = scenes("world8")
    o = ob.OracleScene(w)
    rng = np.random.default_rng(7)
    n = 20000
    org = rng.uniform(-6, 6, (n, 3)).astype(np.float32)
    org[:, 1] = rng.uniform(0, 14, n)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    boxes = o.instance_boxes()
    # aim a third of the rays at cube corners / edge midpoints (silhouette and diagonal cases)
    k = n // 3
    pick = rng.integers(0, len(boxes), k)
    corner = np.where(rng.integers(0, 2, (k, 3)) == 1, boxes[pick, 3:], boxes[pick, :3])
    mid = rng.integers(0, 3, k)
    centre = 0.5 * (boxes[pick, :3] + boxes[pick, 3:])
    corner[np.arange(k), mid] = np.where(rng.integers(0, 2, k) == 1, centre[np.arange(k), mid], corner[np.arange(k), mid])
    d[:k] = corner - org[:k]
    d[k:k + 300] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 300)] * rng.choice([-1, 1], (300, 1))
    d[k + 300:k + 310] = 0
    oi, ot, otime = o.cast_rays(org, d, use_bvh=1)
    gi, gt, gtime = s.cast_rays(org, d, use_bvh=1)
    assert (gi == oi).all() and (gt == ot).all()
    assert (gtime.view(np.uint32) == otime.view(np.uint32)).all()
    assert (oi >= 0).sum() > n // 10


def test_resolution_and_depth_overrides(pkg, gpu_ok):
    """BASELINE configs change width/height/depth through the JSON keys; parity holds there too."""
    w = pkg.World(scene_path("world8_stress"), width=333, height=217, depth=5)
    s = pkg.Scene(w)
    o = ob.OracleScene(w)
    oc, ost = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    c, st = s.render()
    assert (c == oc).all() and st["rays_total"] == ost["rays_total"]
    assert st["waves"] >= 3
    s.close()


def test_tiled_sharding_reassembles(pkg, scenes):
    """world ranks render disjoint tiles; concatenated slabs + rt_untile == the single-GPU frame
    (emulated on one device: ranks run one after another)."""
    import torch
    w, s = scenes("world4")
    full, _ = s.render()
    world = 3
    p0 = s.params(flags=pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_CANVAS_DEVICE, world=world, tile_w=32, tile_h=16)
    slab = s.tiled_slab_pixels(p0)
    gathered = torch.zeros(world * slab, dtype=torch.int32, device="cuda")
    for r in range(world):
        s.render(flags=pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_CANVAS_DEVICE, rank=r, world=world, tile_w=32, tile_h=16,
                 out_ptr=gathered.data_ptr() + 4 * r * slab)
    out = torch.zeros(w.height * w.width, dtype=torch.int32, device="cuda")
    s.untile(p0, gathered.data_ptr(), out.data_ptr())
    got = out.cpu().numpy().view(np.uint32).reshape(w.height, w.width)
    assert (got == full).all()


def test_debug_cast_and_errors(pkg, scenes):
    w, s = scenes("world4")
    o = ob.OracleScene(w)
    inst, tri, t = o.trace_primary(use_bvh=1)
    p = int(np.flatnonzero(inst >= 0)[100])
    x, y = p % w.width, p // w.width
    h = s.debug_cast(x, y)
    assert h.instance == inst[p] and h.triangle == tri[p] and np.float32(h.t) == t[p]
    with pytest.raises(RuntimeError):
        s.debug_cast(-1, 0)
    with pytest.raises(RuntimeError):
        s.render(tile_w=7)
    lib = pkg.cuda_lib()
    assert lib.rt_scene_create(None, None) == 1
    assert b"null" in lib.rt_last_error_string()


def test_empty_scene_and_single_triangle(pkg, gpu_ok):
    """Edge cases: no instances at all; one triangle (LBVH with a single leaf)."""
    C = pkg.C
    w = pkg.World(scene_path("world1"), width=64, height=48)
    d = w.desc
    n_keep = d.n_instances
    d.n_instances = 0
    s = pkg.Scene(w)
    c, st = s.render()
    assert not c.any() and st["rays_total"] == 64 * 48
    s.close()
    d.n_instances = n_keep


@pytest.mark.parametrize("base", ["world4", "world8_stress"])
def test_many_lights_vs_oracle(pkg, gpu_ok, tmp_path, base):
    """Seven lights (the shipped worlds have two): per-light shadow queues, culled lights and the
    bounce waves against the oracle; also a scene without any light."""
    import json
    cfg = json.load(open(scene_path(base)))
    cfg["lights"]["point"] += [{"pos": [x, 6.0, z], "col": [90, 60, 40, 255]} for x, z in ((-3, -2), (3, 1), (0, -4))]
    cfg["lights"]["directional"] += [{"dir": [1.0, -1.0, 0.0], "col": [40, 40, 80, 255]},
                                     {"dir": [0.0, 1.0, 0.0], "col": [30, 30, 30, 255]}]
    p = tmp_path / "lights7.json"
    p.write_text(json.dumps(cfg))
    w = pkg.World(str(p), width=320, height=240)
    s = pkg.Scene(w)
    o = ob.OracleScene(w)
    oc, ost = o.render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    c, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH)
    assert (c == oc).all(), f"{(c != oc).sum()} pixels differ"
    assert st["rays_total"] == ost["rays_total"]
    c2, st2 = s.render(flags=pkg.RT_RENDER_TRACE_UNLIT)
    assert (c2 == oc).all() and st2["rays_culled"] == 0 and st2["rays_total"] == ost["rays_total"]
    s.close()
    cfg["lights"] = {"point": [], "directional": []}
    p.write_text(json.dumps(cfg))
    w = pkg.World(str(p), width=160, height=120)
    s = pkg.Scene(w)
    oc, ost = ob.OracleScene(w).render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    c, st = s.render()
    assert (c == oc).all() and st["rays_total"] == ost["rays_total"]
    s.close()


def test_async_frame_and_external_stream(pkg, scenes):
    """RT_RENDER_ASYNC + rt_frame_finish, on the scene's own stream and on a caller's stream, give
    the frame and the statistics of the blocking call."""
    import torch
    w, s = scenes("world8_stress")
    ref, st_ref = s.render(flags=pkg.RT_RENDER_REBUILD_BVH)
    out = np.zeros((s.height, s.width), np.uint32)
    c, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_ASYNC, out=out)
    assert st is None
    st = s.finish()
    assert (out == ref).all() and st["rays_total"] == st_ref["rays_total"] and st["waves"] == st_ref["waves"]
    with pytest.raises(RuntimeError):
        s.finish()  # nothing pending
    stream = torch.cuda.Stream()
    dev = torch.zeros(s.height * s.width, dtype=torch.int32, device="cuda")
    s.set_stream(stream.cuda_stream)
    try:
        with torch.cuda.stream(stream):
            s.render(flags=pkg.RT_RENDER_ASYNC | pkg.RT_RENDER_CANVAS_DEVICE, out_ptr=dev.data_ptr())
            host = dev.cpu()  # ordered after the frame on the same stream
        st2 = s.finish()
        assert (host.numpy().view(np.uint32).reshape(ref.shape) == ref).all() and st2["rays_total"] == st_ref["rays_total"]
        # a second frame without finish(): the pending one is completed implicitly
        s.render(flags=pkg.RT_RENDER_ASYNC | pkg.RT_RENDER_CANVAS_DEVICE, out_ptr=dev.data_ptr())
        c3, st3 = s.render()
        assert (c3 == ref).all()
    finally:
        s.set_stream(None)


@pytest.mark.parametrize("name", ["world16", "world8_stress"])
def test_chunked_host_frame_equals_single_pass(pkg, scenes, name):
    """A host-canvas frame cut into chunks (chunk c copied to the host while c + 1 renders) is the frame of a
    single pass, pixel for pixel and ray for ray; odd chunk counts and more chunks than tile rows included."""
    w, s = scenes(name)
    oc, ost = ob.OracleScene(w).render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    for chunks, th in ((1, 16), (3, 16), (4, 4), (8, 16), (7, 64)):
        c, st = s.render(flags=pkg.RT_RENDER_STATS, host_chunks=chunks, tile_h=th, tile_w=32)
        assert (c == oc).all(), (chunks, th, int((c != oc).sum()))
        assert st["rays_total"] == ost["rays_total"] and st["rays_secondary"] == ost["rays_secondary"]
        assert st["ms_primary"] > 0 and st["ms_trace"] >= st["ms_primary"]
    s.set_canvas_size(333, 101)  # ragged: width not a multiple of 8, height not a multiple of the tile
    ref, _ = s.render(host_chunks=1)
    got, _ = s.render(host_chunks=5, tile_w=40, tile_h=12)
    assert ref.shape == (101, 333) and (got == ref).all()
    s.set_canvas_size(w.width, w.height)


def test_host_bands_scatter_into_one_canvas(pkg, scenes):
    """RT_RENDER_HOST_BANDS: every rank copies its full-width bands straight into ONE row-major host canvas
    (emulated on one device: the ranks run one after another into the same pinned canvas)."""
    w, s = scenes("world8_stress")
    full, st_full = s.render(host_chunks=1)
    flags = pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_HOST_BANDS
    for world, th, chunks in ((3, 4, 0), (8, 4, 3), (2, 8, 8), (5, 28, 2)):
        canvas, handle = pkg.host_canvas_open(w.width * w.height)
        rays = 0
        for r in range(world):
            _, st = s.render(flags=flags, rank=r, world=world, tile_w=w.width, tile_h=th, out_ptr=canvas.ctypes.data,
                             host_chunks=chunks)
            rays += st["rays_total"]
        assert (canvas.reshape(w.height, w.width) == full).all(), (world, th)
        assert rays == st_full["rays_total"]
        pkg.host_canvas_close(handle)
    # ragged canvas: tile_w > width, last band cut by the canvas edge
    s.set_canvas_size(333, 101)
    ref, _ = s.render(host_chunks=1)
    canvas, handle = pkg.host_canvas_open(333 * 101)
    for r in range(3):
        s.render(flags=flags, rank=r, world=3, tile_w=336, tile_h=8, out_ptr=canvas.ctypes.data, host_chunks=2)
    assert (canvas.reshape(101, 333) == ref).all()
    pkg.host_canvas_close(handle)
    s.set_canvas_size(w.width, w.height)
    # bands need the tiled slab and full-width tiles
    with pytest.raises(RuntimeError):
        s.render(flags=flags, rank=0, world=2, tile_w=32, tile_h=4)


def test_shared_host_canvas_between_processes(pkg, scenes):
    """The shared-memory canvas: a second process that maps it by name sees what this one rendered into it."""
    import subprocess, sys as _sys
    w, s = scenes("world4")
    name = "/rtb200_test_%d" % os.getpid()
    canvas, handle = pkg.host_canvas_open(w.width * w.height, shm_name=name, create=True)
    try:
        full, _ = s.render(host_chunks=1)
        s.render(flags=pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_HOST_BANDS, rank=0, world=1, tile_w=w.width, tile_h=4,
                 out_ptr=canvas.ctypes.data)
        assert (canvas.reshape(w.height, w.width) == full).all()
        code = ("import numpy as np, sys; a = np.memmap('/dev/shm%s', dtype=np.uint32, mode='r'); "
                "print(int(a.astype(np.uint64).sum()))" % name)
        out = subprocess.run([_sys.executable, "-c", code], capture_output=True, text=True, check=True).stdout
        assert int(out.strip()) == int(full.astype(np.uint64).sum())
    finally:
        pkg.host_canvas_close(handle, unlink=True)


def test_kernels_write_a_pinned_host_canvas_directly(pkg, scenes):
    """RT_RENDER_HOST_DIRECT with a page-locked host canvas: no staging canvas and no copy (host_chunks == 0 in the
    statistics), the kernels store the pixels over PCIe; same frame as the staged path and as the oracle. Pageable
    memory falls back to the staged copy."""
    w, s = scenes("world8_stress")
    oc, ost = ob.OracleScene(w).render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    canvas, handle = pkg.host_canvas_open(w.width * w.height)
    D = pkg.RT_RENDER_HOST_DIRECT
    for flags, want_chunks in ((D, 0), (D | pkg.RT_RENDER_FUSED_SHADE, 0), (0, 1), (D | pkg.RT_RENDER_NO_BVH, 0)):
        canvas[:] = 0xdeadbeef
        _, st = s.render(flags=flags, out_ptr=canvas.ctypes.data)
        assert st["host_chunks"] == want_chunks
        if not flags & pkg.RT_RENDER_NO_BVH:
            assert (canvas.reshape(w.height, w.width) == oc).all(), flags
            assert st["rays_total"] == ost["rays_total"]
    # bands of several ranks, written directly into the one canvas
    canvas[:] = 0xdeadbeef
    for r in range(3):
        s.render(flags=D | pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_HOST_BANDS, rank=r, world=3, tile_w=w.width, tile_h=4,
                 out_ptr=canvas.ctypes.data)
    assert (canvas.reshape(w.height, w.width) == oc).all()
    pkg.host_canvas_close(handle)
    pageable = np.empty((w.height, w.width), np.uint32)
    _, st = s.render(flags=D, out=pageable)
    assert st["host_chunks"] == 1 and (pageable == oc).all()


def test_signalled_tile_push(pkg, scenes):
    """rt_push_tiles_signal on one rank: tiles land in the peer canvas, the rank's mailbox slot carries the epoch and
    the wait kernel returns (a missing signal would run into its timeout and leave the epoch unset)."""
    import time
    import torch
    w, s = scenes("world4")
    full, _ = s.render()
    n = w.width * w.height
    ptr, _handle = pkg.peer_canvas_alloc(n)
    flags = pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_CANVAS_DEVICE
    p = s.params(flags=flags, world=1, tile_w=32, tile_h=16)
    slab = torch.zeros(s.tiled_slab_pixels(p), dtype=torch.int32, device="cuda")
    s.render(flags=flags, tile_w=32, tile_h=16, out_ptr=slab.data_ptr())
    t0 = time.time()
    for epoch in (1, 2, 3):
        s.push_tiles_signal(p, slab.data_ptr(), [ptr], n, epoch)
    assert time.time() - t0 < 1.0
    got = s.copy_to_host(ptr, np.empty(n + 16, np.uint32))
    assert (got[:n].reshape(w.height, w.width) == full).all()
    assert got[n] == 3 and (got[n + 1:] == 0).all()
    with pytest.raises(RuntimeError):
        s.push_tiles_signal(p, slab.data_ptr(), [ptr, ptr], n, 4)  # one canvas per rank
    pkg.peer_canvas_free(ptr)


@pytest.mark.parametrize("name,kw", [("world8_stress", {}), ("grid64_mixed", MIXED), ("spheres8", {}), ("world4", {})])
def test_sorted_bounce_waves(pkg, scenes, name, kw):
    """RT_RENDER_SORT_RAYS: per-wave launches with the live paths sorted by origin cell and direction octant between
    the waves; the order of a wave's queue enters no result, so frame and query counts are the oracle's."""
    w, s = scenes(name, **kw)
    oc, ost = ob.OracleScene(w).render(semantics=pkg.RT_SEMANTICS_GPU, use_bvh=1)
    c, st = s.render(flags=pkg.RT_RENDER_SORT_RAYS | pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_STATS)
    assert (c == oc).all(), int((c != oc).sum())
    assert st["rays_total"] == ost["rays_total"] and st["rays_secondary"] == ost["rays_secondary"]
    c2, st2 = s.render()
    assert (c2 == c).all() and st2["waves"] == st["waves"]
