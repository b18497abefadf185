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
