"""gpu-ray-tracer_b200 — Python face of the B200-native ray-traversal engine.

Thin ctypes bindings over two C-ABI shared libraries built in-tree:

* ``librtb200_host.so``  (include/rtb200_host.h) — the C++ scene loader mirroring the reference's
  ``procedural::generate`` / ``SceneBuilder`` (reference: src/procedural/cube_world.cc:38-224);
* ``librtb200.so``       (include/rtb200.h)      — the hand-written sm_100a CUDA engine behind the
  reference's ``rtracer::gpu::update_scene`` seam (reference: include/raytracer.h:18-28).

There is NO CPU fallback: every compute entry point raises ``RuntimeError`` when the CUDA library
or a GPU is missing. The CPU oracle lives under ``oracle/`` and is only reachable from ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s baseline legs — never from here.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)
HOST_LIB_PATH = os.path.join(_HERE, "librtb200_host.so")
CUDA_LIB_PATH = os.path.join(_HERE, "librtb200.so")

RT_SEMANTICS_GPU, RT_SEMANTICS_CPU = 0, 1
RT_RENDER_REBUILD_BVH, RT_RENDER_NO_BVH, RT_RENDER_CANVAS_DEVICE = 1, 2, 4
RT_RENDER_SORT_RAYS, RT_RENDER_TILED_OUTPUT = 8, 16


class RtMaterial(C.Structure):
    _fields_ = [("Ke", C.c_float * 4), ("Ka", C.c_float * 4), ("Kd", C.c_float * 4), ("Ks", C.c_float * 4),
                ("Kt", C.c_float * 4), ("Kr", C.c_float * 4), ("alpha", C.c_float), ("eta", C.c_float)]


class RtTriangle(C.Structure):
    _fields_ = [("v", C.c_int32 * 3), ("tex_degenerate", C.c_int32), ("mat", RtMaterial)]


class RtPose(C.Structure):
    _fields_ = [("rot", C.c_float * 4), ("pos", C.c_float * 3)]


class RtMesh(C.Structure):
    _fields_ = [("pose", RtPose), ("tri_begin", C.c_int32), ("tri_end", C.c_int32)]


class RtInstance(C.Structure):
    _fields_ = [("pose", RtPose), ("mesh", C.c_int32)]


class RtLight(C.Structure):
    _fields_ = [("kind", C.c_int32), ("color", C.c_float * 4), ("v", C.c_float * 3)]


class RtCamera(C.Structure):
    _fields_ = [("pose", RtPose), ("global_near", C.c_float), ("unit_to_pixels", C.c_float),
                ("canvas_width", C.c_float), ("canvas_height", C.c_float)]


class RtEnv(C.Structure):
    _fields_ = [("ambience", C.c_float * 4), ("dist_atten", C.c_float * 3), ("depth", C.c_int32)]


class RtSceneDesc(C.Structure):
    _fields_ = [("vertices", C.POINTER(C.c_float)), ("normals", C.POINTER(C.c_float)), ("n_vertices", C.c_int32),
                ("triangles", C.POINTER(RtTriangle)), ("n_triangles", C.c_int32),
                ("meshes", C.POINTER(RtMesh)), ("n_meshes", C.c_int32),
                ("instances", C.POINTER(RtInstance)), ("n_instances", C.c_int32),
                ("lights", C.POINTER(RtLight)), ("n_lights", C.c_int32),
                ("camera", RtCamera), ("env", RtEnv), ("width", C.c_int32), ("height", C.c_int32)]


class RtRenderParams(C.Structure):
    _fields_ = [("flags", C.c_int32), ("semantics", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
                ("tile_w", C.c_int32), ("tile_h", C.c_int32)]


class RtFrameStats(C.Structure):
    _fields_ = [("rays_total", C.c_uint64), ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("rays_secondary", C.c_uint64), ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("prims_prefiltered", C.c_uint64), ("ms_build", C.c_float), ("ms_trace", C.c_float),
                ("ms_total", C.c_float), ("waves", C.c_int32), ("kernel_launches", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RtBuildStats(C.Structure):
    _fields_ = [("n_prims", C.c_int32), ("n_bvh2_nodes", C.c_int32), ("n_bvh8_nodes", C.c_int32),
                ("kernel_launches", C.c_int32), ("ms_build", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RtHit(C.Structure):
    _fields_ = [("instance", C.c_int32), ("triangle", C.c_int32), ("t", C.c_float), ("normal", C.c_float * 3),
                ("color", C.c_float * 4)]


_host = None
_cuda = None


def host_lib():
    """librtb200_host.so (scene loader; no GPU needed)."""
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise RuntimeError(f"{HOST_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        lib = C.CDLL(HOST_LIB_PATH)
        lib.rth_world_load.argtypes = [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]
        lib.rth_world_load_text.argtypes = lib.rth_world_load.argtypes
        lib.rth_world_desc.argtypes = [C.c_void_p]
        lib.rth_world_desc.restype = C.POINTER(RtSceneDesc)
        lib.rth_world_free.argtypes = [C.c_void_p]
        lib.rth_world_free.restype = None
        lib.rth_last_error_string.restype = C.c_char_p
        lib.rth_world_dump.argtypes = [C.c_void_p, C.c_char_p]
        _host = lib
    return _host


def cuda_lib():
    """librtb200.so (the CUDA engine). Fails loudly when it is not built."""
    global _cuda
    if _cuda is None:
        if not os.path.exists(CUDA_LIB_PATH):
            raise RuntimeError(f"{CUDA_LIB_PATH} is missing: the CUDA engine is not built and there is no CPU "
                               "fallback. Run `python -c 'import __graft_entry__ as g; g.build()'`.")
        lib = C.CDLL(CUDA_LIB_PATH)
        vp, i32, f32p, i32p = C.c_void_p, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int32)
        lib.rt_init.argtypes = [i32]
        lib.rt_device_count.argtypes = [i32p]
        lib.rt_scene_create.argtypes = [C.POINTER(RtSceneDesc), C.POINTER(vp)]
        lib.rt_scene_destroy.argtypes = [vp]
        lib.rt_scene_set_camera.argtypes = [vp, C.POINTER(RtCamera)]
        lib.rt_scene_set_env.argtypes = [vp, C.POINTER(RtEnv)]
        lib.rt_scene_set_instances.argtypes = [vp, C.POINTER(RtInstance), i32]
        lib.rt_scene_set_canvas_size.argtypes = [vp, i32, i32]
        lib.rt_bvh_build.argtypes = [vp, C.POINTER(RtBuildStats)]
        lib.rt_bvh_free.argtypes = [vp]
        lib.rt_update_scene.argtypes = [vp, i32, i32, vp]
        lib.rt_render.argtypes = [vp, C.POINTER(RtRenderParams), vp, C.POINTER(RtFrameStats)]
        lib.rt_tiled_slab_pixels.argtypes = [vp, C.POINTER(RtRenderParams), C.POINTER(C.c_int64)]
        lib.rt_untile.argtypes = [vp, C.POINTER(RtRenderParams), vp, vp]
        lib.rt_trace_primary.argtypes = [vp, i32, vp, vp, vp]
        lib.rt_cast_rays.argtypes = [vp, i32, vp, vp, i32, vp, vp, vp]
        lib.rt_debug_cast.argtypes = [vp, i32, i32, i32, C.POINTER(RtHit)]
        lib.rt_bvh_download.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
        lib.rt_bvh8_download.argtypes = [vp, vp, i32, i32p, vp, i32, i32p]
        lib.rt_measure_l2_bandwidth.argtypes = [C.c_int64, i32, f32p]
        lib.rt_last_error_string.restype = C.c_char_p
        lib.rt_version.restype = C.c_char_p
        _cuda = lib
    return _cuda


def _check(rc, what):
    if rc != 0:
        msg = cuda_lib().rt_last_error_string()
        raise RuntimeError(f"{what} failed with status {rc}: {msg.decode() if msg else ''}")


class World:
    """A loaded world*.json: owns the flat scene arrays (host memory)."""

    def __init__(self, path=None, text=None, width=0, height=0, depth=-1, grid_size=0):
        lib = host_lib()
        handle = C.c_void_p()
        if text is not None:
            rc = lib.rth_world_load_text(text.encode(), width, height, depth, grid_size, C.byref(handle))
        else:
            rc = lib.rth_world_load(os.fspath(path).encode(), width, height, depth, grid_size, C.byref(handle))
        if rc != 0:
            raise RuntimeError("world load failed: " + lib.rth_last_error_string().decode())
        self._h = handle
        self.desc = lib.rth_world_desc(handle).contents
        self.path = path

    @property
    def width(self):
        return self.desc.width

    @property
    def height(self):
        return self.desc.height

    @property
    def n_instances(self):
        return self.desc.n_instances

    def dump(self, path):
        host_lib().rth_world_dump(self._h, os.fspath(path).encode())

    def close(self):
        if self._h:
            host_lib().rth_world_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
