"""GPU: concurrent lanes (rt_scene_set_lanes / RT_FRAME_LANES): a device-resident frame cut into ranges of
screen tiles that run the frame's kernels on streams of their own leaves the canvas, the ray counts and the hit count
of the frame rendered on one stream — and that frame is the oracle's (tests/test_gpu_parity.py). Frames that cannot use
lanes fall back to one stream."""
import numpy as np
import pytest

from conftest import scene_path

pytestmark = pytest.mark.gpu

KEYS = ("rays_total", "rays_primary", "rays_shadow", "rays_culled", "rays_secondary", "n_hits", "waves")


def settle(s, flags, out_ptr, **kw):
    """The wave-0 shadow kernel is chosen from the first six frames of a tile shape; lanes start after that."""
    st = None
    for _ in range(7):
        _, st = s.render(flags=flags, out_ptr=out_ptr, **kw)
    return st


@pytest.mark.parametrize("name,kw,lanes", [("world16", dict(width=1280, height=720, depth=9), 2),
                                           ("world16", dict(width=1920, height=1080, depth=9), 4),
                                           ("world8", dict(width=801, height=603), 3),
                                           ("grid64", dict(width=640, height=480), 8)])
def test_lanes_leave_the_same_frame(pkg, gpu_ok, name, kw, lanes):
    import torch
    w = pkg.World(scene_path(name), **kw)
    s = pkg.Scene(w)
    W, H = w.width, w.height
    base = pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE
    ref = torch.zeros(H * W, dtype=torch.int32, device="cuda")
    dev = torch.zeros(H * W, dtype=torch.int32, device="cuda")
    s.set_lanes(1)
    st0 = settle(s, base, ref.data_ptr())
    assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == 1
    want = ref.cpu().numpy().view(np.uint32)
    host, _ = s.render(flags=pkg.RT_RENDER_REBUILD_BVH)  # the frame the parity tests compare with the oracle
    assert (host.reshape(-1) == want).all()
    s.set_lanes(lanes)
    for i in range(3):
        dev.fill_(0x55)
        _, st = s.render(flags=base, out_ptr=dev.data_ptr())
        assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == lanes, i
        got = dev.cpu().numpy().view(np.uint32)
        assert (got == want).all(), (i, int((got != want).sum()))
        for k in KEYS:
            assert st[k] == st0[k], (i, k, st[k], st0[k])
        assert st["host_chunks"] == 1 and st["ms_trace"] > 0
    # the BVH kept across frames (no rebuild in the frame) and an asynchronous frame
    dev.fill_(0x55)
    s.render(flags=pkg.RT_RENDER_CANVAS_DEVICE | pkg.RT_RENDER_ASYNC, out_ptr=dev.data_ptr())
    st = s.finish()
    assert (dev.cpu().numpy().view(np.uint32) == want).all() and st["rays_total"] == st0["rays_total"]
    # frames that cannot use lanes run on one stream: statistics build, host canvas, graph
    _, st = s.render(flags=base | pkg.RT_RENDER_STATS, out_ptr=dev.data_ptr())
    assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == 1 and st["nodes_visited"] > 0
    assert (dev.cpu().numpy().view(np.uint32) == want).all()
    host, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH)
    assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == 1 and (host.reshape(-1) == want).all()
    for i in range(3):
        dev.fill_(0x55)
        s.render(flags=base | pkg.RT_RENDER_GRAPH, out_ptr=dev.data_ptr())
        assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == 1
        assert (dev.cpu().numpy().view(np.uint32) == want).all(), i
    # and lanes again after them
    dev.fill_(0x55)
    s.render(flags=base, out_ptr=dev.data_ptr())
    assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == lanes
    assert (dev.cpu().numpy().view(np.uint32) == want).all()
    with pytest.raises(RuntimeError):
        s.set_lanes(9)
    s.close()


@pytest.mark.parametrize("world,lanes", [(2, 2), (8, 3)])
def test_lanes_on_a_rank_of_a_sharded_frame(pkg, gpu_ok, world, lanes):
    """Rank r of `world` (tile slab output, the multi-GPU frame of bench.py --gather nccl / p2p): same slab with lanes."""
    import torch
    w = pkg.World(scene_path("world16"), width=1920, height=1080, depth=9)
    s = pkg.Scene(w)
    flags = pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE | pkg.RT_RENDER_TILED_OUTPUT
    for rank in (0, world - 1):
        p = s.params(flags=flags, rank=rank, world=world, tile_w=32, tile_h=16)
        n = s.tiled_slab_pixels(p)
        ref = torch.zeros(n, dtype=torch.int32, device="cuda")
        dev = torch.zeros(n, dtype=torch.int32, device="cuda")
        s.set_lanes(1)
        st0 = settle(s, flags, ref.data_ptr(), rank=rank, world=world, tile_w=32, tile_h=16)
        s.set_lanes(lanes)
        for i in range(2):
            _, st = s.render(flags=flags, rank=rank, world=world, tile_w=32, tile_h=16, out_ptr=dev.data_ptr())
            assert s.debug_get(pkg.RT_DEBUG_FRAME_LANES) == lanes
            # the slab padding beyond the canvas edge is never written: compare what a rank owns through the tile map
            idx, slab = pkg.tile_map(w.width, w.height, 32, 16, rank, world)
            own = idx[idx >= 0]
            a, b = dev.cpu().numpy().view(np.uint32), ref.cpu().numpy().view(np.uint32)
            assert slab == n and (a[own] == b[own]).all()
            for k in KEYS:
                assert st[k] == st0[k], (rank, k)
    s.close()
