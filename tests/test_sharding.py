"""CPU: the multi-GPU path's host logic. Screen tiles are dealt round-robin to ranks, every rank
fills a compact slab, the slabs are all-gathered and de-interleaved. Here two gloo ranks fill their
slabs on the host with the pixel ids the tile map assigns them (the map is the code the kernels use,
rt_tile_map) and the gathered result must be the identity image."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.environ["RT_ROOT"])
import rtb200_loader
pkg = rtb200_loader.load()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
W, H, tw, th = 203, 97, 32, 16        # ragged right and bottom tiles
idx, slab = pkg.tile_map(W, H, tw, th, rank, world)
mine = idx >= 0
local = torch.full((slab,), -7, dtype=torch.int64)
pix = np.arange(W * H).reshape(H, W)
local[torch.from_numpy(idx[mine])] = torch.from_numpy(pix[mine])      # "render": pixel id as colour
gathered = torch.empty(world * slab, dtype=torch.int64)
dist.all_gather_into_tensor(gathered, local)
out = np.full((H, W), -1, np.int64)
owners = np.zeros((H, W), int)
for r in range(world):                                                # de-interleave (k_untile on the GPU)
    ridx, _ = pkg.tile_map(W, H, tw, th, r, world)
    m = ridx >= 0
    owners += m
    out[m] = gathered[r * slab:(r + 1) * slab].numpy()[ridx[m]]
assert (owners == 1).all(), "every pixel has exactly one owner"
assert (out == pix).all(), "gathered canvas is the identity image"
# balance: round-robin tiles differ by at most one tile between ranks
counts = [int((pkg.tile_map(W, H, tw, th, r, world)[0] >= 0).sum()) for r in range(world)]
assert max(counts) - min(counts) <= tw * th
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_two_rank_gloo_gather(tmp_path, pkg):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, RT_ROOT=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29517", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_tile_map_partitions_the_canvas(pkg, world):
    W, H = 640, 480
    seen = np.zeros((H, W), int)
    for r in range(world):
        idx, slab = pkg.tile_map(W, H, 32, 16, r, world)
        m = idx >= 0
        seen += m
        assert idx[m].max() < slab and len(np.unique(idx[m])) == m.sum()
    assert (seen == 1).all()


def test_tile_map_rejects_bad_tiles(pkg):
    with pytest.raises(RuntimeError):
        pkg.tile_map(64, 64, 7, 16)
    with pytest.raises(RuntimeError):
        pkg.tile_map(64, 64, 32, 16, rank=2, world=2)


BANDS_WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.environ["RT_ROOT"])
import rtb200_loader
pkg = rtb200_loader.load()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
W, H, th = 203, 97, 4                 # width not a multiple of 8, last band cut by the canvas edge
tw = (W + 7) // 8 * 8                 # full-width bands (RT_RENDER_HOST_BANDS wants tile_w >= width)
canvas = np.memmap(os.environ["RT_SHM"], dtype=np.int64, mode="r+", shape=(H, W))
idx, slab = pkg.tile_map(W, H, tw, th, rank, world)
local = np.full(slab, -7, np.int64)
pix = np.arange(W * H).reshape(H, W)
mine = idx >= 0
local[idx[mine]] = pix[mine]          # "render" the rank's bands into its slab
# the copy rt_render does for a banded frame (copy_rows_to_host): band k of this rank = slab rows [k th, (k + 1) th) with
# pitch tile_w -> canvas rows (k world + rank) th ..., cut at the canvas edge
n_bands = (H + th - 1) // th
k = 0
while k * world + rank < n_bands:
    y0 = (k * world + rank) * th
    rows = min(th, H - y0)
    canvas[y0:y0 + rows, :] = local[k * th * tw:(k * th + rows) * tw].reshape(rows, tw)[:, :W]
    k += 1
canvas.flush()
dist.barrier()
if rank == 0:
    assert (np.asarray(canvas) == pix).all(), "the ranks' bands make up the identity image in the shared canvas"
rows_owned = int(mine.any(axis=1).sum())
t = torch.tensor([rows_owned])
dist.all_reduce(t)
assert int(t.item()) == H
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


@pytest.mark.parametrize("world", [2, 3])
def test_host_bands_fill_one_shared_canvas(tmp_path, pkg, world):
    """The host-frame path of N > 1 (RT_RENDER_HOST_BANDS) without a GPU: gloo ranks fill their band slabs through the
    kernels' tile map and copy them into ONE canvas in shared memory with the row arithmetic of copy_rows_to_host."""
    script = tmp_path / "bands_worker.py"
    script.write_text(BANDS_WORKER)
    shm = "/dev/shm/rtb200_cpu_bands_%d_%d" % (os.getpid(), world)
    np.memmap(shm, dtype=np.int64, mode="w+", shape=(97, 203))[:] = -1
    try:
        env = dict(os.environ, RT_ROOT=ROOT, RT_SHM=shm, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29520 + world), WORLD_SIZE=str(world))
        procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                                  stderr=subprocess.STDOUT, text=True) for r in range(world)]
        outs = [p.communicate(timeout=300)[0] for p in procs]
        for p, o in zip(procs, outs):
            assert p.returncode == 0, o
            assert "ok" in o
    finally:
        os.unlink(shm)
