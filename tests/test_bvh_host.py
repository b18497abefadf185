"""CPU: structural invariants of the host LBVH / BVH8 reference build (oracle/lbvh_host.c) that the
GPU build is compared against bit for bit in tests/test_gpu_parity.py."""
import numpy as np
import pytest

import oracle_binding as ob
from conftest import scene_path


def decode_children(node):
    """-> list of (slot, kind, lo[3], hi[3], payload) for the 8 slots of one 80-byte node."""
    f = node.view(np.float32)
    org = f[0:3]
    e = [(int(node[3]) >> (8 * k)) & 0xff for k in range(3)]
    imask = int(node[3]) >> 24
    scale = [np.float32(2.0) ** (x - 127) for x in e]
    b = node[4:].view(np.uint8)  # child_base(4) prim_base(4) meta(8) qlo xyz (24) qhi xyz (24)
    meta = b[8:16]
    q = b[16:].reshape(6, 8)
    out = []
    for s in range(8):
        if meta[s] == 0:
            continue
        lo = np.array([org[k] + np.float32(q[k][s]) * scale[k] for k in range(3)], np.float32)
        hi = np.array([org[k] + np.float32(q[3 + k][s]) * scale[k] for k in range(3)], np.float32)
        inner = (meta[s] & 0x18) == 0x18
        assert inner == bool(imask >> s & 1)
        out.append((s, "inner" if inner else "leaf", lo, hi, int(meta[s])))
    return out, int(node[4]), int(node[5]), imask


@pytest.mark.parametrize("name", ["world1", "world4", "world16"])
def test_bvh8_covers_every_triangle_once_and_boxes_are_conservative(pkg, name):
    w = pkg.World(scene_path(name))
    b = ob.OracleBvh(w)
    n = b.n_prims
    assert n == 12 * w.n_instances
    assert sorted(b.perm.tolist()) == list(range(n))
    assert sorted(b.prim_order8.tolist()) == list(range(n))
    assert (np.diff(b.morton.astype(np.uint64)) >= 0).all()
    seen = np.zeros(n, int)
    visited = 0
    stack = [(0, None, None)]
    while stack:
        idx, plo, phi = stack.pop()
        visited += 1
        kids, child_base, prim_base, imask = decode_children(b.nodes8[idx])
        rank = 0
        for s, kind, lo, hi, meta in kids:
            assert (hi >= lo).all()
            if kind == "inner":
                stack.append((child_base + rank, lo, hi))
                rank += 1
            else:
                cnt = {1: 1, 3: 2, 7: 3}[meta >> 5]
                off = meta & 31
                for j in range(cnt):
                    p = prim_base + off + j
                    seen[p] += 1
                    tri = b.prims8[p]
                    v0, e1, e2 = tri[0:3], tri[4:7], tri[8:11]
                    for v in (v0, v0 + e1, v0 + e2):
                        assert (v >= lo).all() and (v <= hi).all()
    assert visited == b.n_nodes8
    assert (seen == 1).all()


def test_degenerate_sizes(pkg):
    """0, 1, 2 and 3 triangles: trees without internal BVH2 nodes / with a single leaf slot."""
    w = pkg.World(scene_path("world1"))
    d = w.desc
    keep_inst, keep_end = d.n_instances, d.meshes[0].tri_end
    try:
        d.n_instances = 0
        assert ob.OracleBvh(w).n_prims == 0
        d.n_instances = 1
        for k in (1, 2, 3, 5):
            d.meshes[d.instances[0].mesh].tri_end = d.meshes[d.instances[0].mesh].tri_begin + k
            b = ob.OracleBvh(w)
            assert b.n_prims == k and b.n_nodes8 >= 1
            assert sorted(b.prim_order8.tolist()) == list(range(k))
    finally:
        d.n_instances = keep_inst
        d.meshes[0].tri_end = keep_end
