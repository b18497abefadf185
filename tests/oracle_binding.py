"""ctypes binding of the CPU oracle (oracle/librt_oracle.so). TEST INFRASTRUCTURE: importable from
tests/, __graft_entry__.smoke() and bench.py's baseline legs only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "librt_oracle.so")


class OrStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("rays_total", "rays_primary", "rays_shadow", "rays_secondary",
                                           "box_tests", "leaf_tests", "tri_tests")]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "rt_oracle.c")
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "librt_oracle.so"], stdout=subprocess.DEVNULL)
        l = C.CDLL(LIB)
        vp, i32 = C.c_void_p, C.c_int32
        l.or_scene_create.argtypes = [vp]
        l.or_scene_create.restype = vp
        l.or_scene_destroy.argtypes = [vp]
        l.or_scene_set_camera.argtypes = [vp, vp]
        l.or_scene_set_env.argtypes = [vp, vp]
        l.or_scene_set_canvas_size.argtypes = [vp, i32, i32]
        l.or_instance_boxes.argtypes = [vp, vp]
        l.or_render.argtypes = [vp, i32, i32, i32, i32, vp, C.POINTER(OrStats), i32]
        l.or_trace_primary.argtypes = [vp, i32, vp, vp, vp, i32]
        l.or_cast_rays.argtypes = [vp, i32, vp, vp, i32, vp, vp, vp]
        l.or_debug_cast.argtypes = [vp, i32, i32, i32, i32, vp]
        l.or_primary_ray.argtypes = [vp, i32, i32, vp, vp]
        l.or_canvas_hash.argtypes = [vp, C.c_int64, C.POINTER(i32)]
        l.or_canvas_hash.restype = C.c_uint64
        _lib = l
    return _lib


def canvas_hash(canvas):
    canvas = np.ascontiguousarray(canvas, dtype=np.uint32)
    nz = C.c_int32()
    h = lib().or_canvas_hash(canvas.ctypes.data, canvas.size, C.byref(nz))
    return h, nz.value


class OracleScene:
    def __init__(self, world):
        self.world = world  # keeps the arrays alive
        self._h = lib().or_scene_create(C.addressof(world.desc))
        self.width, self.height = world.width, world.height
        self.n_instances = world.n_instances

    def set_canvas_size(self, w, h):
        lib().or_scene_set_canvas_size(self._h, w, h)
        self.width, self.height = w, h

    def set_camera(self, cam):
        lib().or_scene_set_camera(self._h, C.addressof(cam))

    def set_env(self, env):
        lib().or_scene_set_env(self._h, C.addressof(env))

    def render(self, semantics=0, use_bvh=1, threads=None, rows=None):
        canvas = np.zeros((self.height, self.width), dtype=np.uint32)
        st = OrStats()
        y0, y1 = rows if rows else (0, self.height)
        lib().or_render(self._h, semantics, use_bvh, y0, y1, canvas.ctypes.data, C.byref(st),
                        threads or os.cpu_count() or 1)
        return canvas, st.as_dict()

    def trace_primary(self, use_bvh=1, threads=None):
        n = self.width * self.height
        inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
        lib().or_trace_primary(self._h, use_bvh, inst.ctypes.data, tri.ctypes.data, t.ctypes.data,
                               threads or os.cpu_count() or 1)
        return inst, tri, t

    def cast_rays(self, origins, dirs, use_bvh=1):
        o = np.ascontiguousarray(origins, np.float32); d = np.ascontiguousarray(dirs, np.float32)
        n = o.shape[0]
        inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
        lib().or_cast_rays(self._h, use_bvh, o.ctypes.data, d.ctypes.data, n, inst.ctypes.data, tri.ctypes.data,
                           t.ctypes.data)
        return inst, tri, t

    def instance_boxes(self):
        b = np.empty((self.n_instances, 6), np.float32)
        lib().or_instance_boxes(self._h, b.ctypes.data)
        return b

    def primary_ray(self, x, y):
        o = np.empty(3, np.float32); d = np.empty(3, np.float32)
        lib().or_primary_ray(self._h, x, y, o.ctypes.data, d.ctypes.data)
        return o, d

    def close(self):
        if self._h:
            lib().or_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- host LBVH / BVH8 reference build (oracle/lbvh_host.c) ----
class OracleBvh:
    def __init__(self, world):
        l = lib()
        vp = C.c_void_p
        l.or_bvh_build.argtypes = [vp]
        l.or_bvh_build.restype = vp
        l.or_bvh_free.argtypes = [vp]
        for name in ("n_prims", "n_nodes8", "levels"):
            f = getattr(l, "or_bvh_" + name)
            f.argtypes = [vp]
            f.restype = C.c_int32
        for name in ("morton", "perm", "left", "right", "parent", "node_boxes", "bounds", "nodes8", "prims8", "prim_order8"):
            f = getattr(l, "or_bvh_" + name)
            f.argtypes = [vp]
            f.restype = vp
        self.world = world
        self._h = l.or_bvh_build(C.addressof(world.desc))
        n = self.n_prims = l.or_bvh_n_prims(self._h)
        self.n_nodes8 = l.or_bvh_n_nodes8(self._h)
        self.levels = l.or_bvh_levels(self._h)

        def arr(name, dtype, count):
            if count <= 0:
                return np.zeros(0, dtype)
            p = getattr(l, "or_bvh_" + name)(self._h)
            return np.ctypeslib.as_array(C.cast(p, C.POINTER(np.ctypeslib.as_ctypes_type(dtype))), shape=(count,)).copy()

        self.morton = arr("morton", np.uint64, n)
        self.perm = arr("perm", np.int32, n)
        self.left = arr("left", np.int32, max(n - 1, 0))
        self.right = arr("right", np.int32, max(n - 1, 0))
        self.parent = arr("parent", np.int32, max(2 * n - 1, 0))
        self.node_boxes = arr("node_boxes", np.float32, 6 * max(2 * n - 1, 0)).reshape(-1, 6)
        self.bounds = arr("bounds", np.float32, 6)
        self.nodes8 = arr("nodes8", np.uint32, 20 * self.n_nodes8).reshape(-1, 20)
        self.prims8 = arr("prims8", np.float32, 12 * n).reshape(-1, 12)
        self.prim_order8 = arr("prim_order8", np.int32, n)

    def close(self):
        if self._h:
            lib().or_bvh_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
