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
# RTB200_CUDA_LIB: an experiment build of the engine (tools/build_variant.sh) instead of the shipped library
CUDA_LIB_PATH = os.environ.get("RTB200_CUDA_LIB") or os.path.join(_HERE, "librtb200.so")

RT_SEMANTICS_GPU, RT_SEMANTICS_CPU = 0, 1
RT_RENDER_REBUILD_BVH, RT_RENDER_NO_BVH, RT_RENDER_CANVAS_DEVICE = 1, 2, 4
RT_RENDER_SORT_RAYS, RT_RENDER_TILED_OUTPUT, RT_RENDER_STATS, RT_RENDER_FUSED_SHADE = 8, 16, 32, 64
RT_RENDER_TRACE_UNLIT = 128
RT_RENDER_ASYNC = 256
RT_RENDER_REFIT_BVH = 512
RT_RENDER_HOST_BANDS = 1024
RT_RENDER_HOST_DIRECT = 2048
RT_RENDER_SPLIT_SHADE = 4096
RT_RENDER_PEER_STORE = 8192
RT_RENDER_TILE_KERNEL = 16384
RT_RENDER_GRAPH = 32768
RT_RENDER_FAST_ACCEPT = 65536
RT_HITABLE_TRIMESH, RT_HITABLE_SPHERE = 0, 1
RT_DEBUG_PREFIX_UPLOADS, RT_DEBUG_BVH8_LEVELS, RT_DEBUG_BUILD_PATH, RT_DEBUG_BUILD_LAUNCHES, RT_DEBUG_BUILD_CLUSTER = 0, 1, 2, 3, 4
RT_DEBUG_GRAPH_REPLAYS, RT_DEBUG_GRAPH_CAPTURES, RT_DEBUG_FRAME_LANES = 5, 6, 7
RT_DEBUG_BUILD_PHASE_NS = 16
RT_DEBUG_BUILD_LEVEL_NS = 32
RT_CAST_STATIC_MAX_RAYS = 65536


class RtMaterial(C.Structure):
    _fields_ = [("Ke", C.c_float * 4), ("Ka", C.c_float * 4), ("Kd", C.c_float * 4), ("Ks", C.c_float * 4),
                ("Kt", C.c_float * 4), ("Kr", C.c_float * 4), ("alpha", C.c_float), ("eta", C.c_float)]


class RtTriangle(C.Structure):
    _fields_ = [("v", C.c_int32 * 3), ("tex_degenerate", C.c_int32), ("mat", RtMaterial),
                ("tex_x", C.c_int32), ("tex_y", C.c_int32), ("tex_u", C.c_float * 2), ("tex_v", C.c_float * 2)]


class RtPose(C.Structure):
    _fields_ = [("rot", C.c_float * 4), ("pos", C.c_float * 3)]


class RtMesh(C.Structure):
    _fields_ = [("pose", RtPose), ("tri_begin", C.c_int32), ("tri_end", C.c_int32), ("kind", C.c_int32), ("radius", C.c_float)]


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
                ("camera", RtCamera), ("env", RtEnv), ("width", C.c_int32), ("height", C.c_int32),
                ("atlas", C.POINTER(C.c_float)), ("atlas_width", C.c_int32), ("atlas_height", C.c_int32)]


class RtRenderParams(C.Structure):
    _fields_ = [("flags", C.c_int32), ("semantics", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
                ("tile_w", C.c_int32), ("tile_h", C.c_int32), ("host_chunks", C.c_int32)]


class RtFrameStats(C.Structure):
    _fields_ = [("rays_total", C.c_uint64), ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("rays_secondary", C.c_uint64), ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("prims_prefiltered", C.c_uint64), ("ms_build", C.c_float), ("ms_trace", C.c_float),
                ("ms_total", C.c_float), ("waves", C.c_int32), ("kernel_launches", C.c_int32),
                ("ms_primary", C.c_float), ("ms_shade", C.c_float), ("ms_bounce", C.c_float), ("n_hits", C.c_int32), ("ms_shadow", C.c_float),
                ("ms_wave_phase", C.c_float * 4), ("rays_culled", C.c_uint64), ("host_chunks", C.c_int32),
                ("tile_kernel", C.c_int32), ("node_fetches", C.c_uint64), ("graph", C.c_int32)]

    def as_dict(self):
        return {k: (list(getattr(self, k)) if k == "ms_wave_phase" else getattr(self, k)) for k, _ in self._fields_}


class RtBuildStats(C.Structure):
    _fields_ = [("n_prims", C.c_int32), ("n_bvh2_nodes", C.c_int32), ("n_bvh8_nodes", C.c_int32),
                ("kernel_launches", C.c_int32), ("ms_build", C.c_float), ("n_levels8", C.c_int32)]

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
        f3, f4 = C.POINTER(C.c_float * 3), C.POINTER(C.c_float * 4)
        lib.rth_pose_translate_global.argtypes = [C.POINTER(RtPose), f3]
        lib.rth_pose_translate.argtypes = [C.POINTER(RtPose), f3]
        lib.rth_pose_rotate.argtypes = [C.POINTER(RtPose), f4]
        lib.rth_quat_axis_angle.argtypes = [f3, C.c_float, f4]
        lib.rth_camera_key.argtypes = [C.POINTER(RtCamera), C.c_char]
        lib.rth_camera_mouse.argtypes = [C.POINTER(RtCamera), C.c_float, C.c_float]
        for fn in (lib.rth_pose_translate_global, lib.rth_pose_translate, lib.rth_pose_rotate, lib.rth_quat_axis_angle,
                   lib.rth_camera_mouse):
            fn.restype = None
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
        lib.rt_bvh_refit.argtypes = [vp, C.POINTER(RtBuildStats)]
        lib.rt_update_scene.argtypes = [vp, i32, i32, vp]
        lib.rt_render.argtypes = [vp, C.POINTER(RtRenderParams), vp, C.POINTER(RtFrameStats)]
        lib.rt_tiled_slab_pixels.argtypes = [vp, C.POINTER(RtRenderParams), C.POINTER(C.c_int64)]
        lib.rt_untile.argtypes = [vp, C.POINTER(RtRenderParams), vp, vp]
        lib.rt_scene_set_stream.argtypes = [vp, vp, i32]
        lib.rt_scene_set_lanes.argtypes = [vp, i32]
        lib.rt_peer_canvas_alloc.argtypes = [C.c_int64, C.POINTER(vp), vp]
        lib.rt_peer_canvas_free.argtypes = [vp]
        lib.rt_peer_canvas_open.argtypes = [vp, C.POINTER(vp)]
        lib.rt_peer_canvas_close.argtypes = [vp]
        lib.rt_push_tiles.argtypes = [vp, C.POINTER(RtRenderParams), vp, C.POINTER(vp), i32]
        lib.rt_push_tiles_signal.argtypes = [vp, C.POINTER(RtRenderParams), vp, C.POINTER(vp), i32, C.c_int64, C.c_uint32]
        lib.rt_copy_to_host.argtypes = [vp, vp, vp, C.c_int64]
        lib.rt_scene_set_peers.argtypes = [vp, C.POINTER(vp), i32, C.c_int64, i32]
        lib.rt_peer_frame_canvas.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(C.c_uint32)]
        lib.rt_host_canvas_open.argtypes = [C.c_char_p, C.c_int64, i32, C.POINTER(vp)]
        lib.rt_host_canvas_close.argtypes = [vp, C.c_int64, C.c_char_p, i32]
        lib.rt_frame_finish.argtypes = [vp, C.POINTER(RtFrameStats)]
        lib.rt_tile_map.argtypes = [i32, i32, i32, i32, i32, i32, vp, C.POINTER(C.c_int64)]
        lib.rt_trace_primary.argtypes = [vp, i32, vp, vp, vp]
        lib.rt_cast_rays.argtypes = [vp, i32, vp, vp, i32, vp, vp, vp]
        lib.rt_cast_rays_static.argtypes = [vp, i32, vp, vp, i32, vp, vp, vp]
        lib.rt_debug_cast.argtypes = [vp, i32, i32, i32, C.POINTER(RtHit)]
        lib.rt_bvh_download.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
        lib.rt_bvh8_download.argtypes = [vp, vp, i32, i32p, vp, vp, i32, i32p]
        lib.rt_debug_get.argtypes = [vp, i32, C.POINTER(C.c_int64)]
        lib.rt_measure_l2_bandwidth.argtypes = [C.c_int64, i32, f32p]
        lib.rt_last_error_string.restype = C.c_char_p
        lib.rt_version.restype = C.c_char_p
        _cuda = lib
    return _cuda


def _check(rc, what):
    if rc != 0:
        msg = cuda_lib().rt_last_error_string()
        raise RuntimeError(f"{what} failed with status {rc}: {msg.decode() if msg else ''}")


def peer_canvas_alloc(pixels):
    """(device pointer, 64-byte CUDA IPC handle) of a canvas other ranks can map."""
    ptr = C.c_void_p()
    handle = (C.c_ubyte * 64)()
    _check(cuda_lib().rt_peer_canvas_alloc(pixels, C.byref(ptr), handle), "rt_peer_canvas_alloc")
    return ptr.value, bytes(handle)


def peer_canvas_open(handle):
    ptr = C.c_void_p()
    buf = (C.c_ubyte * 64).from_buffer_copy(handle)
    _check(cuda_lib().rt_peer_canvas_open(buf, C.byref(ptr)), "rt_peer_canvas_open")
    return ptr.value


def host_canvas_open(pixels, shm_name=None, create=False):
    """Page-locked host canvas as a uint32 numpy array: private (shm_name None) or POSIX shared memory that every
    rank of the box maps (RT_RENDER_HOST_BANDS). Keep the returned handle for host_canvas_close()."""
    ptr = C.c_void_p()
    name = shm_name.encode() if shm_name else None
    _check(cuda_lib().rt_host_canvas_open(name, pixels, 1 if create else 0, C.byref(ptr)), "rt_host_canvas_open")
    arr = np.ctypeslib.as_array(C.cast(ptr.value, C.POINTER(C.c_uint32)), shape=(pixels,))
    return arr, (ptr.value, pixels, name)


def host_canvas_close(handle, unlink=False):
    ptr, pixels, name = handle
    _check(cuda_lib().rt_host_canvas_close(ptr, pixels, name, 1 if unlink else 0), "rt_host_canvas_close")


def peer_canvas_close(ptr):
    _check(cuda_lib().rt_peer_canvas_close(ptr), "rt_peer_canvas_close")


def peer_canvas_free(ptr):
    _check(cuda_lib().rt_peer_canvas_free(ptr), "rt_peer_canvas_free")


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

    def apply_moves(self, spec):
        """Per-frame motion in the reference's arithmetic (include/rtb200_host.h), in place on this world's
        camera and instances. Comma-separated operations, the grammar of oracle/ref_build/ref_driver.cu
        --moves: ``w s a d`` (camera keys, src/main.cc:140-165), ``m:xrel:yrel`` (mouse look, main.cc:166-177),
        ``g:idx:dx:dy:dz`` / ``l:idx:dx:dy:dz`` (Entity::translate_global / translate of instance idx),
        ``r:idx:ax:ay:az:theta`` (Entity::rotate by Quat(axis, theta)). Pass the result on with
        Scene.set_camera(world.desc.camera) / Scene.set_instances(world.desc.instances, n)."""
        lib = host_lib()
        cam = self.desc.camera
        for op in filter(None, spec.split(",")):
            f = op.split(":")
            if f[0] in "wsad" and len(f) == 1:
                lib.rth_camera_key(C.byref(cam), f[0].encode())
            elif f[0] == "m" and len(f) == 3:
                lib.rth_camera_mouse(C.byref(cam), float(f[1]), float(f[2]))
            elif f[0] in ("g", "l") and len(f) == 5:
                pose = self.desc.instances[int(f[1])].pose
                dp = (C.c_float * 3)(float(f[2]), float(f[3]), float(f[4]))
                (lib.rth_pose_translate_global if f[0] == "g" else lib.rth_pose_translate)(C.byref(pose), C.byref(dp))
            elif f[0] == "r" and len(f) == 6:
                pose = self.desc.instances[int(f[1])].pose
                axis = (C.c_float * 3)(float(f[2]), float(f[3]), float(f[4]))
                q = (C.c_float * 4)()
                lib.rth_quat_axis_angle(C.byref(axis), float(f[5]), C.byref(q))
                lib.rth_pose_rotate(C.byref(pose), C.byref(q))
            else:
                raise ValueError("bad move operation %r" % op)

    def close(self):
        if self._h:
            host_lib().rth_world_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Scene:
    """A scene resident on one GPU. Mirrors the reference's per-frame entry points:
    ``update_scene`` (rtracer::gpu::update_scene, src/raytracer.cu:102-120) and ``debug_cast``
    (src/raytracer.cu:91-100), plus the full-control ``render`` and the parity hooks."""

    def __init__(self, world, device=0):
        lib = cuda_lib()
        _check(lib.rt_init(device), "rt_init")
        self.world = world
        self.device = device
        h = C.c_void_p()
        _check(lib.rt_scene_create(C.byref(world.desc), C.byref(h)), "rt_scene_create")
        self._h = h
        self.width, self.height = world.width, world.height

    # -- host writes the reference does on its managed Scene
    def set_camera(self, cam):
        _check(cuda_lib().rt_scene_set_camera(self._h, C.byref(cam)), "rt_scene_set_camera")

    def set_env(self, env):
        _check(cuda_lib().rt_scene_set_env(self._h, C.byref(env)), "rt_scene_set_env")

    def set_instances(self, instances, n):
        _check(cuda_lib().rt_scene_set_instances(self._h, instances, n), "rt_scene_set_instances")

    def set_canvas_size(self, w, h):
        _check(cuda_lib().rt_scene_set_canvas_size(self._h, w, h), "rt_scene_set_canvas_size")
        self.width, self.height = w, h

    # -- BVH
    def build_bvh(self):
        st = RtBuildStats()
        _check(cuda_lib().rt_bvh_build(self._h, C.byref(st)), "rt_bvh_build")
        return st.as_dict()

    def refit_bvh(self):
        """Keeps the tree, recomputes every box after instances moved (rt_bvh_refit)."""
        st = RtBuildStats()
        _check(cuda_lib().rt_bvh_refit(self._h, C.byref(st)), "rt_bvh_refit")
        return st.as_dict()

    def free_bvh(self):
        _check(cuda_lib().rt_bvh_free(self._h), "rt_bvh_free")

    # -- frames
    def update_scene(self, kernel_dim=16, optimize=True, out=None):
        canvas = out if out is not None else np.empty((self.height, self.width), np.uint32)
        _check(cuda_lib().rt_update_scene(self._h, kernel_dim, 1 if optimize else 0, canvas.ctypes.data), "rt_update_scene")
        return canvas

    def params(self, flags=0, semantics=RT_SEMANTICS_GPU, rank=0, world=1, tile_w=0, tile_h=0, host_chunks=0):
        return RtRenderParams(flags, semantics, rank, world, tile_w, tile_h, host_chunks)

    def render(self, flags=0, semantics=RT_SEMANTICS_GPU, rank=0, world=1, tile_w=0, tile_h=0, out=None, out_ptr=None,
               host_chunks=0):
        """Returns (canvas, stats). out: a host uint32 array to fill; out_ptr: a raw pointer (host, or
        device with RT_RENDER_CANVAS_DEVICE) — then canvas is None."""
        p = self.params(flags, semantics, rank, world, tile_w, tile_h, host_chunks)
        st = RtFrameStats()
        canvas = None
        if out_ptr is None:
            if flags & RT_RENDER_HOST_BANDS:
                canvas = out if out is not None else np.zeros((self.height, self.width), np.uint32)
            elif flags & RT_RENDER_TILED_OUTPUT:
                canvas = out if out is not None else np.empty(self.tiled_slab_pixels(p), np.uint32)
            else:
                canvas = out if out is not None else np.empty((self.height, self.width), np.uint32)
            out_ptr = canvas.ctypes.data
        _check(cuda_lib().rt_render(self._h, C.byref(p), out_ptr, C.byref(st)), "rt_render")
        return canvas, (None if flags & RT_RENDER_ASYNC else st.as_dict())

    def set_stream(self, cuda_stream):
        """Enqueue all later work of this scene on the caller's CUDA stream (an int handle, e.g.
        torch.cuda.current_stream().cuda_stream); None returns to the scene's own stream."""
        _check(cuda_lib().rt_scene_set_stream(self._h, cuda_stream or None, 0 if cuda_stream is None else 1), "rt_scene_set_stream")

    def set_lanes(self, lanes):
        """rt_scene_set_lanes: concurrent lanes of a device-resident frame (0 = library default, RT_FRAME_LANES or 1)."""
        _check(cuda_lib().rt_scene_set_lanes(self._h, lanes), "rt_scene_set_lanes")

    def finish(self):
        """Waits for the frame started with RT_RENDER_ASYNC and returns its statistics."""
        st = RtFrameStats()
        _check(cuda_lib().rt_frame_finish(self._h, C.byref(st)), "rt_frame_finish")
        return st.as_dict()

    def tiled_slab_pixels(self, params):
        n = C.c_int64()
        _check(cuda_lib().rt_tiled_slab_pixels(self._h, C.byref(params), C.byref(n)), "rt_tiled_slab_pixels")
        return n.value

    def untile(self, params, gathered_dev_ptr, canvas_dev_ptr):
        _check(cuda_lib().rt_untile(self._h, C.byref(params), gathered_dev_ptr, canvas_dev_ptr), "rt_untile")

    def push_tiles(self, params, slab_dev_ptr, peer_ptrs):
        """Stores this rank's tiles into every rank's row-major canvas (NVLink peer memory)."""
        arr = (C.c_void_p * len(peer_ptrs))(*peer_ptrs)
        _check(cuda_lib().rt_push_tiles(self._h, C.byref(params), slab_dev_ptr, arr, len(peer_ptrs)), "rt_push_tiles")

    def push_tiles_signal(self, params, slab_dev_ptr, peer_ptrs, canvas_pixels, epoch):
        """push_tiles + the exchange barrier over the canvases' mailboxes (no collective call)."""
        arr = (C.c_void_p * len(peer_ptrs))(*peer_ptrs)
        _check(cuda_lib().rt_push_tiles_signal(self._h, C.byref(params), slab_dev_ptr, arr, len(peer_ptrs), canvas_pixels, epoch),
               "rt_push_tiles_signal")

    def set_peers(self, canvases, n_ranks, canvas_pixels, gather_root=-1):
        """canvases: 2 * n_ranks device pointers, [parity][rank] (rt_scene_set_peers); frames rendered with
        RT_RENDER_PEER_STORE are stored into them by the kernels that produce the pixels."""
        arr = (C.c_void_p * len(canvases))(*canvases)
        _check(cuda_lib().rt_scene_set_peers(self._h, arr, n_ranks, canvas_pixels, gather_root), "rt_scene_set_peers")

    def peer_frame_canvas(self, rank):
        """(device pointer, epoch) of the canvas of `rank` that holds the last RT_RENDER_PEER_STORE frame."""
        ptr, ep = C.c_void_p(), C.c_uint32()
        _check(cuda_lib().rt_peer_frame_canvas(self._h, rank, C.byref(ptr), C.byref(ep)), "rt_peer_frame_canvas")
        return ptr.value, ep.value

    def copy_to_host(self, dev_ptr, out):
        _check(cuda_lib().rt_copy_to_host(self._h, dev_ptr, out.ctypes.data, out.size), "rt_copy_to_host")
        return out

    # -- parity hooks
    def trace_primary(self, use_bvh=1):
        n = self.width * self.height
        inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
        _check(cuda_lib().rt_trace_primary(self._h, use_bvh, inst.ctypes.data, tri.ctypes.data, t.ctypes.data),
               "rt_trace_primary")
        return inst, tri, t

    def cast_rays(self, origins, dirs, use_bvh=1):
        o = np.ascontiguousarray(origins, np.float32); d = np.ascontiguousarray(dirs, np.float32)
        n = o.shape[0]
        inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
        _check(cuda_lib().rt_cast_rays(self._h, use_bvh, o.ctypes.data, d.ctypes.data, n, inst.ctypes.data,
                                       tri.ctypes.data, t.ctypes.data), "rt_cast_rays")
        return inst, tri, t

    def cast_rays_static(self, origins, dirs, use_bvh=1):
        """rt_cast_rays_static: rt_cast_rays over the library's static device scratch (no allocation, at most
        RT_CAST_STATIC_MAX_RAYS rays, an existing BVH)."""
        o = np.ascontiguousarray(origins, np.float32); d = np.ascontiguousarray(dirs, np.float32)
        n = o.shape[0]
        inst = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
        _check(cuda_lib().rt_cast_rays_static(self._h, use_bvh, o.ctypes.data, d.ctypes.data, n, inst.ctypes.data,
                                              tri.ctypes.data, t.ctypes.data), "rt_cast_rays_static")
        return inst, tri, t

    def debug_cast(self, x, y, semantics=RT_SEMANTICS_GPU):
        h = RtHit()
        _check(cuda_lib().rt_debug_cast(self._h, x, y, semantics, C.byref(h)), "rt_debug_cast")
        return h

    def download_bvh(self):
        lib = cuda_lib()
        nn, npr = C.c_int32(), C.c_int32()
        _check(lib.rt_bvh8_download(self._h, None, 0, C.byref(nn), None, None, 0, C.byref(npr)), "rt_bvh8_download")
        n, n8 = npr.value, nn.value
        out = dict(n_prims=n, n_nodes8=n8)
        out["morton"] = np.zeros(n, np.uint64); out["perm"] = np.zeros(n, np.int32)
        out["left"] = np.zeros(max(n - 1, 0), np.int32); out["right"] = np.zeros(max(n - 1, 0), np.int32)
        out["parent"] = np.zeros(max(2 * n - 1, 0), np.int32)
        out["node_boxes"] = np.zeros((max(2 * n - 1, 0), 6), np.float32); out["bounds"] = np.zeros(6, np.float32)
        _check(lib.rt_bvh_download(self._h, out["morton"].ctypes.data, out["perm"].ctypes.data, out["left"].ctypes.data,
                                   out["right"].ctypes.data, out["parent"].ctypes.data, out["node_boxes"].ctypes.data,
                                   out["bounds"].ctypes.data), "rt_bvh_download")
        out["nodes8"] = np.zeros((n8, 20), np.uint32); out["prims8"] = np.zeros((n, 12), np.float32)
        out["prim_order8"] = np.zeros(n, np.int32)
        _check(lib.rt_bvh8_download(self._h, out["nodes8"].ctypes.data, n8, C.byref(nn), out["prims8"].ctypes.data,
                                    out["prim_order8"].ctypes.data, n, C.byref(npr)), "rt_bvh8_download")
        return out

    def debug_get(self, what):
        v = C.c_int64()
        _check(cuda_lib().rt_debug_get(self._h, what, C.byref(v)), "rt_debug_get")
        return v.value

    def close(self):
        if getattr(self, "_h", None):
            cuda_lib().rt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def measure_l2_bandwidth(nbytes=32 << 20, iters=50):
    g = C.c_float()
    _check(cuda_lib().rt_measure_l2_bandwidth(nbytes, iters, C.byref(g)), "rt_measure_l2_bandwidth")
    return g.value


def tile_map(width, height, tile_w=32, tile_h=16, rank=0, world=1):
    """Host-only: (slab index per pixel or -1, slab size) — the tile map the kernels use."""
    idx = np.empty(width * height, np.int64)
    n = C.c_int64()
    _check(cuda_lib().rt_tile_map(width, height, tile_w, tile_h, rank, world, idx.ctypes.data, C.byref(n)), "rt_tile_map")
    return idx.reshape(height, width), n.value
