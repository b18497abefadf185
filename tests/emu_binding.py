"""ctypes binding of tests/emu/libemu.so: the engine's __host__ __device__ traversal / shading
code compiled for the CPU (TEST INFRASTRUCTURE; see tests/emu/emu.cu)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "emu", "libemu.so")
_lib = None


def available():
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(LIB)
        vp, i32 = C.c_void_p, C.c_int32
        l.emu_trace_primary.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp, i32]
        l.emu_cast_rays.argtypes = [vp, vp, vp, i32, i32, vp, vp, i32, vp, vp, vp]
        l.emu_render.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp, i32]
        l.emu_check_identity_rotation.argtypes = [vp, i32]
        # emu_warp.cu: rt_beam.cuh on an emulated warp
        l.emu_beam_primary.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp, i32]
        l.emu_beam_shadow.argtypes = [vp, vp, vp, i32, i32, vp, vp, i32]
        l.emu_beam_cast_rays.argtypes = [vp, vp, vp, i32, vp, vp, i32, vp, vp, vp]
        _lib = l
    return _lib


def trace_primary(world, bvh, mode, threads=None):
    n = world.width * world.height
    inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
    st = np.zeros(3, np.uint64)
    nodes = np.ascontiguousarray(bvh.nodes8); prims = np.ascontiguousarray(bvh.prims8)
    lib().emu_trace_primary(C.addressof(world.desc), nodes.ctypes.data, prims.ctypes.data, bvh.n_prims, mode,
                            inst.ctypes.data, tri.ctypes.data, t.ctypes.data, st.ctypes.data, threads or os.cpu_count())
    return inst, tri, t, dict(nodes=int(st[0]), prims=int(st[1]), exact=int(st[2]))


def cast_rays(world, bvh, mode, origins, dirs):
    o = np.ascontiguousarray(origins, np.float32); d = np.ascontiguousarray(dirs, np.float32)
    n = o.shape[0]
    inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
    nodes = np.ascontiguousarray(bvh.nodes8); prims = np.ascontiguousarray(bvh.prims8)
    lib().emu_cast_rays(C.addressof(world.desc), nodes.ctypes.data, prims.ctypes.data, bvh.n_prims, mode,
                        o.ctypes.data, d.ctypes.data, n, inst.ctypes.data, tri.ctypes.data, t.ctypes.data)
    return inst, tri, t


def render(world, bvh, mode, semantics, threads=None):
    canvas = np.zeros((world.height, world.width), np.uint32)
    rays = np.zeros(3, np.uint64)
    nodes = np.ascontiguousarray(bvh.nodes8); prims = np.ascontiguousarray(bvh.prims8)
    lib().emu_render(C.addressof(world.desc), nodes.ctypes.data, prims.ctypes.data, bvh.n_prims, mode, semantics,
                     canvas.ctypes.data, rays.ctypes.data, threads or os.cpu_count())
    return canvas, dict(primary=int(rays[0]), shadow=int(rays[1]), secondary=int(rays[2]))


def beam_primary(world, bvh, use_box=1, threads=None):
    """Camera rays through rt_beam.cuh's beam_query on emulated warps (one per 8x4 tile), as k_tile<1> calls it."""
    n = world.width * world.height
    inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
    st = np.zeros(4, np.uint64)
    nodes = np.ascontiguousarray(bvh.nodes8); prims = np.ascontiguousarray(bvh.prims8)
    lib().emu_beam_primary(C.addressof(world.desc), nodes.ctypes.data, prims.ctypes.data, bvh.n_prims, use_box,
                           inst.ctypes.data, tri.ctypes.data, t.ctypes.data, st.ctypes.data, threads or os.cpu_count())
    return inst, tri, t, dict(nodes=int(st[0]), prims=int(st[1]), exact=int(st[2]), node_fetches=int(st[3]))


def beam_shadow(world, bvh, cull=0, threads=None):
    """Occlusion of the camera hits' shadow rays per light: (warp-coherent walk, per-lane walk), each n_lights x H x W,
    255 where there is no shadow ray."""
    n_l = max(world.desc.n_lights, 1)
    a = np.empty((n_l, world.height, world.width), np.uint8); b = np.empty_like(a)
    nodes = np.ascontiguousarray(bvh.nodes8); prims = np.ascontiguousarray(bvh.prims8)
    lib().emu_beam_shadow(C.addressof(world.desc), nodes.ctypes.data, prims.ctypes.data, bvh.n_prims, cull,
                          a.ctypes.data, b.ctypes.data, threads or os.cpu_count())
    return a, b


def beam_cast_rays(world, bvh, origins, dirs):
    o = np.ascontiguousarray(origins, np.float32); d = np.ascontiguousarray(dirs, np.float32)
    n = o.shape[0]
    inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
    nodes = np.ascontiguousarray(bvh.nodes8); prims = np.ascontiguousarray(bvh.prims8)
    lib().emu_beam_cast_rays(C.addressof(world.desc), nodes.ctypes.data, prims.ctypes.data, bvh.n_prims,
                             o.ctypes.data, d.ctypes.data, n, inst.ctypes.data, tri.ctypes.data, t.ctypes.data)
    return inst, tri, t
