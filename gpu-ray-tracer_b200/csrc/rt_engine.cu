// rt_engine.cu — scene upload, frame driver and render kernels of librtb200.so.
//
// Replaces, behind the C ABI of include/rtb200.h:
//   SceneBuilder::build_gpu_scene upload half     src/scene_builder.cu:83-179   -> rt_scene_create
//   rtracer::gpu::update_scene / create_bvh       src/raytracer.cu:76-120       -> rt_update_scene / rt_render
//   rtracer::gpu::trace (megakernel)              src/raytracer.cu:17-43        -> k_primary + k_shade + k_bounce
//   rtracer::gpu::debug / debug_cast              src/raytracer.cu:45-52,91-100 -> rt_debug_cast
//   gputils (managed memory, blocking copies)     include/gputils/alloc.h       -> DevBuf, one stream, async copies
//
// Frame = wavefront pipeline on one stream:
//   k_primary  persistent warps pull 8x4-pixel tiles from a counter, generate camera rays
//              (Camera::at), find the closest hit through the BVH8, write black for misses and
//              append hits to a queue with one warp-aggregated atomic (ballot + popc + shuffle).
//   k_shade    one thread per queued hit: normal, Phong, shadow rays through the same BVH8;
//              pixels whose material spawns no secondary ray are finished here, the others get a
//              path slot (RayFrame stack in HBM) and enter the bounce queue.
//   k_bounce   one wave per launch: cast the top frame's ray of every live path, shade, run the
//              reference's REFLECT/REFRACT state machine, compact survivors into the next queue.
// There is no CPU fallback anywhere in this library.
#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "rt_beam.cuh"
#include "rt_bvh.cuh"
#include "rt_hostprep.h"
#include "rt_internal.h"
#include "rt_shade.cuh"

using namespace rtd;

// ---------------------------------------------------------------- errors, buffers
namespace rth {
static thread_local std::string g_err;
unsigned long long g_alloc_gen = 0;
void set_error(const std::string& msg) { g_err = msg; }
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    g_err = std::string(cudaGetErrorName(e)) + ": " + cudaGetErrorString(e) + " in " + what + " (" + file + ":" +
            std::to_string(line) + ")";
    if (e == cudaErrorMemoryAllocation) return RT_ERR_OUT_OF_MEMORY;
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) return RT_ERR_NO_DEVICE;
    return RT_ERR_CUDA;
}
int DevBuf::reserve(size_t bytes) {
    if (bytes <= cap && p) return 0;
    if (bytes == 0) bytes = 16;
    release();
    g_alloc_gen++;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) { p = nullptr; cap = 0; return cuda_fail(e, "cudaMalloc", __FILE__, __LINE__); }
    cap = bytes;
    return 0;
}
void DevBuf::release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}
void Bvh::release() {
    DevBuf* all[] = {&prim_world, &prim_box, &scene_bounds, &keys_a, &keys_b, &vals_a, &vals_b, &hist, &left, &right,
                     &parent, &range_lo, &range_hi, &flags, &node_box, &nodes8, &prims8, &prim_order8, &work,
                     &collapse_scratch, &edge_box, &rec, &node8_root, &node8_slot};
    for (DevBuf* b : all) b->release();
    if (h_off) { cudaFreeHost(h_off); h_off = nullptr; h_off_cap = 0; }
    off_in_flight = false;
    valid = false;
    topology_valid = false;
}
}  // namespace rth
using rth::set_error;

namespace {

constexpr int kThreads = 256;
#ifndef RT_SHADOW_MIN_BLOCKS
#define RT_SHADOW_MIN_BLOCKS 3
#endif
#ifndef RT_WAVES_MIN_BLOCKS
#define RT_WAVES_MIN_BLOCKS 3
#endif
#ifndef RT_RESOLVE_MIN_BLOCKS
#define RT_RESOLVE_MIN_BLOCKS 4   /* 64 registers, no spills: shading phase 0.905 -> 0.884 ms against 3 blocks */
#endif
#ifndef RT_GEN_MIN_BLOCKS
#define RT_GEN_MIN_BLOCKS 4
#endif
static int g_device = -1;
static int g_sm_count = 148;

// ---------------------------------------------------------------- screen tiling
// Work unit = one 8x4-pixel warp tile. Screen tiles of tile_w x tile_h pixels are dealt round-robin
// to `world` ranks; a rank enumerates the warp tiles of the screen tiles it owns.
struct TileMap {
    int W, H, tile_w, tile_h, tiles_x, n_tiles, rank, world, n_owned, sub_x, sub_per_tile, tiled_out;
    int k_begin, k_count;  // the owned tiles [k_begin, k_begin + k_count) this launch works on (a chunk of the frame)
    int reverse;           // k_tile: work units are taken from the last one down (see there)
    __host__ __device__ int n_work() const { return k_count * sub_per_tile; }
    // lane -> pixel; returns false for pixels outside the canvas
    __host__ __device__ bool locate(int q, int lane, int& x, int& y, int& out) const {
        int k = q / sub_per_tile, sidx = q - k * sub_per_tile;
        k += k_begin;
        int T = k * world + rank;
        int tx = T % tiles_x, ty = T / tiles_x;
        int sx = sidx % sub_x, sy = sidx / sub_x;
        int lx = sx * 8 + (lane & 7), ly = sy * 4 + (lane >> 3);
        x = tx * tile_w + lx;
        y = ty * tile_h + ly;
        out = tiled_out ? (k * tile_h + ly) * tile_w + lx : y * W + x;
        return x < W && y < H;
    }
};

static int make_tilemap(const RtScene* sc, const RtRenderParams* p, TileMap& m) {
    int tw = 32, th = 16, rank = 0, world = 1, tiled = 0;
    if (p) {
        if (p->tile_w > 0) tw = p->tile_w;
        if (p->tile_h > 0) th = p->tile_h;
        if (p->world > 1) { world = p->world; rank = p->rank; }
        tiled = (p->flags & RT_RENDER_TILED_OUTPUT) ? 1 : 0;
    }
    if (tw % 8 || th % 4 || tw <= 0 || th <= 0 || rank < 0 || rank >= world) {
        set_error("tile_w must be a multiple of 8, tile_h a multiple of 4, 0 <= rank < world");
        return RT_ERR_INVALID_ARG;
    }
    m.W = sc->width; m.H = sc->height; m.tile_w = tw; m.tile_h = th;
    m.tiles_x = (m.W + tw - 1) / tw;
    m.n_tiles = m.tiles_x * ((m.H + th - 1) / th);
    m.rank = rank; m.world = world;
    m.n_owned = m.n_tiles > rank ? (m.n_tiles - rank + world - 1) / world : 0;
    m.sub_x = tw / 8;
    m.sub_per_tile = m.sub_x * (th / 4);
    m.tiled_out = tiled;
    m.k_begin = 0; m.k_count = m.n_owned;
    static const int order_env = [] { const char* e = std::getenv("RT_TILE_ORDER"); return e ? std::atoi(e) : 0; }();
    m.reverse = order_env;
    return RT_OK;
}
static long long slab_pixels(const TileMap& m) {
    long long per_rank = (m.n_tiles + m.world - 1) / m.world;
    return per_rank * m.tile_w * m.tile_h;
}

// ---------------------------------------------------------------- device counters
// d_counters layout (unsigned long long[32]):
enum { C_WORK0 = 0, C_HITQ = 1, C_WORK1 = 2, C_PATHS = 3, C_QUEUE_A = 4, C_QUEUE_B = 5, C_WORK2 = 6,
       C_RAYS_PRIMARY = 8, C_RAYS_SHADOW = 9, C_RAYS_SECONDARY = 10, C_NODES = 11, C_PRIMS = 12, C_EXACT = 13, C_RAYS_CULLED = 14, C_WAVES = 15, C_BARRIER = 16, C_PHASE_NS = 20, C_FETCHES = 24, C_N = 32 };

__device__ __forceinline__ void flush_counts(unsigned long long* ctr, const RayCounts& rc, const TravStats& ts, bool stats) {
    unsigned v[6] = {rc.primary, rc.shadow, rc.secondary, ts.nodes, ts.prims, ts.exact};
    int n = stats ? 6 : 3;
    for (int i = 0; i < n; i++) {
        unsigned x = v[i];
        for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if ((threadIdx.x & 31) == 0 && x) atomicAdd(&ctr[C_RAYS_PRIMARY + i], (unsigned long long)x);
    }
}

// One atomicAdd per converged group of lanes instead of one per lane (same-address atomics
// serialise in one L2 slice: 2.6 M of them cost k_resolve 0.25 ms on world8_stress at 4K).
__device__ __forceinline__ int warp_alloc(unsigned long long* counter) {
    unsigned m = __activemask();
    int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(m));
    base = __shfl_sync(m, base, leader);
    return (int)base + __popc(m & ((1u << lane) - 1u));
}

__device__ __forceinline__ unsigned warp_alloc32(unsigned* counter) {
    unsigned m = __activemask();
    int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned)__popc(m));
    base = __shfl_sync(m, base, leader);
    return base + __popc(m & ((1u << lane) - 1u));
}

struct HitRec {  // 32 bytes
    int out_index, xy, inst, tri;
    float t, u, v, t_local;
};

// ---------------------------------------------------------------- instance boxes (create_boxes, src/raytracer.cu:54-74)
// BoundingBox::from_local (src/rayopt/bounding_box.cu:51-60) refits only the transformed min and
// max corners — reproduced as is, because the reference culls instances by exactly this box.
__global__ void k_instance_boxes(InstRec* insts, int n, const MeshRec* meshes, const float* mesh_box) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    InstRec I = insts[i];
    instance_box(I, meshes, mesh_box);
    for (int k = 0; k < 6; k++) insts[i].box[k] = I.box[k];
    insts[i].box_ok = I.box_ok;
}

// Scene set-up: the ray-independent part of every mesh triangle's exact test (rt_exact.cuh ref_triangle_pre).
__global__ void k_tri_precompute(const TriRec* tris, int n, const float4* verts, int n_verts, float4* tri_pre) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const TriRec T = tris[i];
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (T.v0 >= 0 && T.v1 >= 0 && T.v2 >= 0 && T.v0 < n_verts && T.v1 < n_verts && T.v2 < n_verts) {  // a sphere's record has no vertices
        const float4 a = verts[T.v0], b = verts[T.v1], c = verts[T.v2];
        const rtx::TriPre p = rtx::ref_triangle_pre(rtx::mk3(a.x, a.y, a.z), rtx::mk3(b.x, b.y, b.z), rtx::mk3(c.x, c.y, c.z));
        o = make_float4(p.n.x, p.n.y, p.n.z, p.area);
    }
    tri_pre[i] = o;
}

// ---------------------------------------------------------------- tracer selection
enum { TR_BVH_BOX = 0, TR_BRUTE = 1, TR_BVH_NOBOX = 2, TR_BRUTE_BOX = 3, TR_BVH_BOX_FAST = 4 };  // _FAST: BVH_BOX in the opt-in tolerance mode
template <int MODE> struct TracerOf;
template <> struct TracerOf<TR_BVH_BOX> {
    typedef BvhTracer T;
    __device__ static T make(const SceneView* s, TravStats* st) { T t; t.s = s; t.use_box = true; t.st = st; return t; }
};
template <> struct TracerOf<TR_BVH_BOX_FAST> {
    typedef BvhTracerT<true> T;
    __device__ static T make(const SceneView* s, TravStats* st) { T t; t.s = s; t.use_box = true; t.st = st; return t; }
};
template <> struct TracerOf<TR_BVH_NOBOX> {
    typedef BvhTracer T;
    __device__ static T make(const SceneView* s, TravStats* st) { T t; t.s = s; t.use_box = false; t.st = st; return t; }
};
template <> struct TracerOf<TR_BRUTE> {
    typedef BruteTracer T;
    __device__ static T make(const SceneView* s, TravStats*) { T t; t.s = s; t.use_box = false; return t; }
};
template <> struct TracerOf<TR_BRUTE_BOX> {
    typedef BruteTracer T;
    __device__ static T make(const SceneView* s, TravStats*) { T t; t.s = s; t.use_box = true; return t; }
};

// ---------------------------------------------------------------- wave 0: primary rays
// IDS == true: parity hook (rt_trace_primary) — writes (instance, triangle-in-mesh, t) per pixel.
template <int MODE, bool STATS, bool IDS>
__global__ void __launch_bounds__(kThreads) k_primary(SceneView s, TileMap tm, unsigned long long* ctr, HitRec* hitq,
                                                     uint32_t* canvas, int* id_inst, int* id_tri, float* id_t, int aligned) {
    TravStats ts = {0, 0, 0};
    RayCounts rc = {0, 0, 0};
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, STATS ? &ts : nullptr);
    const int lane = threadIdx.x & 31;
    const int n_work = tm.n_work();
    for (;;) {
        int q = 0;
        if (lane == 0) q = (int)atomicAdd(&ctr[C_WORK0], 1ull);
        q = __shfl_sync(0xffffffffu, q, 0);
        if (q >= n_work) break;
        int x, y, out;
        bool valid = tm.locate(q, lane, x, y, out);
        Hit h;
        hit_reset(h);
        bool hit = false;
        if (valid) {
            rtx::Ray r = rtx::camera_ray(s.cam, (float)x, (float)y);
            rc.primary++;
            hit = tr.closest(r, h);
        }
        if (IDS) {
            if (valid) {
                int p = y * tm.W + x;
                id_inst[p] = hit ? h.inst : -1;
                id_tri[p] = hit ? h.tri - s.meshes[s.insts[h.inst].mesh].tri_begin : -1;
                id_t[p] = h.t;
            }
            continue;
        }
        if (valid && !hit) canvas[out] = 0u;  // pack_color(0,0,0,0)
        // Warp-aggregated append of the hits. ctr[C_HITQ]: low word = queue slots, high word = hits.
        // aligned: a tile with any hit takes 32 slots, lane l writes slot base + l (inst = -1 for its misses), so that
        // every warp of the shading kernels works on the 8x4 pixels of ONE tile. Packed (popc(m) slots per tile), a
        // warp of k_shadow held the hits of two or three tiles from unrelated parts of the screen — whichever warps
        // of k_primary appended one after the other — and walked the union of their traversals.
        unsigned m = __ballot_sync(0xffffffffu, hit);
        if (m) {
            unsigned long long base = 0;
            const unsigned long long slots = aligned ? 32ull : (unsigned long long)__popc(m);
            if (lane == 0) base = atomicAdd(&ctr[C_HITQ], slots | ((unsigned long long)__popc(m) << 32));
            base = __shfl_sync(0xffffffffu, base, 0) & 0xffffffffull;
            if (hit || aligned) {
                HitRec rec;
                rec.out_index = out; rec.xy = x | (y << 16); rec.inst = hit ? h.inst : -1; rec.tri = h.tri;
                rec.t = h.t; rec.u = h.u; rec.v = h.v; rec.t_local = h.t_local;
                hitq[base + (aligned ? (unsigned)lane : (unsigned)__popc(m & ((1u << lane) - 1u)))] = rec;
            }
        }
    }
    flush_counts(ctr, rc, ts, STATS);
}

// ---------------------------------------------------------------- wave 0: shading of primary hits
template <int MODE, bool STATS, bool OPAQUE_ONLY>
__global__ void __launch_bounds__(kThreads, 3) k_shade(SceneView s, unsigned long long* ctr, const HitRec* hitq, uint32_t* canvas,
                                                   unsigned char* paths, int path_stride, int* queue) {
    TravStats ts = {0, 0, 0};
    RayCounts rc = {0, 0, 0};
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, STATS ? &ts : nullptr);
    const int n = (int)ctr[C_HITQ];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        HitRec rec = hitq[i];
        if (rec.inst < 0) continue;  // a missed pixel inside a tile-aligned group
        Hit h;
        h.t = rec.t; h.t_local = rec.t_local; h.u = rec.u; h.v = rec.v; h.inst = rec.inst; h.tri = rec.tri;
        rtx::Ray r = rtx::camera_ray(s.cam, (float)(rec.xy & 0xffff), (float)((unsigned)rec.xy >> 16));
        int mat_id = s.tris[h.tri].mat;
        const MatRec& m = s.mats[mat_id];
        if (!OPAQUE_ONLY && s.depth > 0 && (m.reflective || m.refractive)) {
            // the pixel continues: give it a path slot; k_bounce shades this hit as its first step
            int slot = warp_alloc(&ctr[C_PATHS]);
            PathHead* head = (PathHead*)(paths + (size_t)slot * path_stride);
            Frame* frames = (Frame*)(head + 1);
            PathHead hd;
            hd.acc = rtx::mk4(0.f, 0.f, 0.f, 0.f);
            hd.out_index = rec.out_index; hd.stack_top = 0; hd.isect_mat = mat_id; hd.pending = 1;
            hd.hit = h; hd.pad0 = hd.pad1 = 0;
            *head = hd;
            Frame f;
            frame_init_root(f, r, s.depth);
            frames[0] = f;
            int qi = warp_alloc(&ctr[C_QUEUE_A]);
            queue[qi] = slot;
            continue;
        }
        // no secondary rays: acc = 0 + 1 * illuminate, clamp, pack (src/raytracer.cu:36-41)
        float3 nrm = hit_normal(s, h, r);
        float4 c = illuminate<OPAQUE_ONLY>(s, tr, r, h.t, nrm, m, diffuse_colour(s, m, h.tri, h.u, h.v), rc);
        canvas[rec.out_index] = rtx::pack_color(c);
    }
    flush_counts(ctr, rc, ts, STATS);
}

// ---------------------------------------------------------------- split shading (every wave)
// The fused kernels (k_shade / k_bounce below, RT_RENDER_FUSED_SHADE) keep ray setup, traversal,
// the exact tests, Phong and the bounce state machine in one instruction stream (50-80 KB of
// SASS, stall_no_instruction 2.6 warps per issue). By default the same work runs as tight kernels:
//   k_shadow_gen   one shadow ray per (hit, light)                    no traversal
//   k_shadow       any-hit traversal, or Light::attenuate's closest-hit loop when a material refracts
//   k_resolve      Phong + pack (wave 0; starts a path where the material reflects / refracts)
//   k_bounce_waves (waves >= 1, one cooperative launch) per wave: closest hit of each live path's
//                  next ray / shadow-ray generation / shadow traversal / Phong + RayFrame state machine
struct ShadowRec { float4 o_maxt, d; };  // 32 bytes; d.w = hit index (int bits). Light l's rays: shq[l * n .. l * n + shq_count[l])

__device__ __forceinline__ Frame* path_frames(unsigned char* paths, int path_stride, int slot, PathHead** head) {
    *head = (PathHead*)(paths + (size_t)slot * path_stride);
    return (Frame*)(*head + 1);
}

// Lane assignment for the traversal phases when a launch has fewer rays than threads (the late bounce
// waves of a deep frame carry a few thousand rays): rays are dealt to every `stride`-th thread of the
// grid instead of to consecutive lanes, stride = the largest power of two <= 32 that still gives every
// ray a thread. The rays of such a wave are the divergent ones (paths caught between reflective
// faces); packed 32 to a warp on a handful of SMs, one warp executed the union of 32 unrelated
// traversals and exact tests (81 us per wave for ~5 k rays on world8_stress at depth 9), spread out
// each warp follows one or two rays and every SM takes part. Returns false for threads without work.
__device__ __forceinline__ bool sparse_lanes(long long n, long long& first, long long& step) {
    const long long T = (long long)gridDim.x * blockDim.x;
    const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int stride = 1;
    while (stride < 32 && n * (stride * 2) <= T) stride <<= 1;
    first = gt / stride;
    step = T / stride;
    return (gt & (stride - 1)) == 0;
}

// PATHS == false: hits of camera rays (wave 0). PATHS == true: hits of path rays; rec.out_index is
// the path slot and rec.inst < 0 marks a miss (no shadow rays).
template <bool PATHS>
__device__ __forceinline__ void phase_shadow_gen(const SceneView& s, const int n, const HitRec* hitq, ShadowRec* shq,
                                                 unsigned* shq_count, unsigned char* occl, float4* att4,
                                                 unsigned char* paths, int path_stride, unsigned& culled) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        HitRec rec = hitq[i];
        rtx::Ray r;
        const bool hit = rec.inst >= 0;  // wave 0: a missed pixel inside a tile-aligned group; waves: a path ray that missed
        if (PATHS) {
            PathHead* head;
            Frame* frames = path_frames(paths, path_stride, rec.out_index, &head);
            if (hit) r = frames[head->stack_top].ray;
        } else if (hit) {
            r = rtx::camera_ray(s.cam, (float)(rec.xy & 0xffff), (float)((unsigned)rec.xy >> 16));
        } else {
            // wave 0 queues are slot-indexed: k_shadow must find a hole, not stale data
            ShadowRec hole;
            hole.o_maxt = make_float4(0.f, 0.f, 0.f, -1.f);
            hole.d = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int l = 0; l < s.n_lights; l++) shq[(size_t)l * n + i] = hole;
        }
        float3 hit_pos = hit ? rtx::ray_at(r, rec.t) : rtx::mk3(0.f, 0.f, 0.f);
        float3 nrm = rtx::mk3(0.f, 0.f, 0.f);
        const MatRec* m = nullptr;
        if (hit && s.cull_unlit) {
            Hit h;
            h.t = rec.t; h.t_local = rec.t_local; h.u = rec.u; h.v = rec.v; h.inst = rec.inst; h.tri = rec.tri;
            nrm = hit_normal(s, h, r);
            m = &s.mats[s.tris[rec.tri].mat];
        }
        if (!hit) continue;
        for (int l = 0; l < s.n_lights; l++) {
            size_t j = (size_t)l * n + i;  // where the resolve step looks the answer up
            float3 dir;
            float max_t, da;
            light_setup(s, s.lights[l], hit_pos, dir, max_t, da);
            if (m && light_unlit(*m, nrm, dir)) {
                // never read by the resolve step (same predicate there); defined anyway
                culled++;
                if (att4) att4[j] = make_float4(0.f, 0.f, 0.f, 0.f); else occl[j] = 1;
                if (!PATHS) {
                    ShadowRec hole;
                    hole.o_maxt = make_float4(0.f, 0.f, 0.f, -1.f);
                    hole.d = make_float4(0.f, 0.f, 0.f, 0.f);
                    shq[j] = hole;
                }
                continue;
            }
            rtx::Ray cur = shadow_ray(hit_pos, dir);
            ShadowRec q;
            q.o_maxt = make_float4(cur.o.x, cur.o.y, cur.o.z, max_t);
            q.d = make_float4(cur.d.x, cur.d.y, cur.d.z, __int_as_float((int)i));
            // PATHS: dense per-light queues (about half of the path rays miss; the holes left k_shadow
            // 7-12 active lanes per warp). Wave 0: slot i, so neighbouring lanes keep neighbouring
            // pixels and no atomics are needed; culled lights leave a hole (max_t < 0).
            shq[(size_t)l * n + (PATHS ? warp_alloc32(&shq_count[l]) : (unsigned)i)] = q;
        }
    }
    if (!PATHS && blockIdx.x == 0 && threadIdx.x < s.n_lights)
        for (int l = threadIdx.x; l < s.n_lights; l += blockDim.x) shq_count[l] = (unsigned)n;
}

__device__ __forceinline__ void flush_culled(const SceneView& s, unsigned long long* ctr, unsigned culled) {
    if (s.cull_unlit) {
        for (int o = 16; o; o >>= 1) culled += __shfl_xor_sync(0xffffffffu, culled, o);
        if ((threadIdx.x & 31) == 0 && culled) atomicAdd(&ctr[C_RAYS_CULLED], (unsigned long long)culled);
    }
}

template <bool PATHS>
__global__ void __launch_bounds__(kThreads, RT_GEN_MIN_BLOCKS) k_shadow_gen(SceneView s, unsigned long long* ctr, int ctr_n,
                                                        const HitRec* hitq, ShadowRec* shq, unsigned* shq_count,
                                                        unsigned char* occl, float4* att4, unsigned char* paths,
                                                        int path_stride) {
    unsigned culled = 0;
    phase_shadow_gen<PATHS>(s, (int)(unsigned)ctr[ctr_n], hitq, shq, shq_count, occl, att4, paths, path_stride, culled);
    flush_culled(s, ctr, culled);
}

// LOOP == false: occl[j] = 1 when anything blocks the ray within max_t (any-hit).
// LOOP == true : att4[j] = Light::attenuate's result (closest-hit walk through refractive occluders).
// work: null = static assignment (entry f to thread f mod T, or sparse lanes for small waves); otherwise a counter
// from which every warp pulls the next 32 entries — shadow rays cost between one node test and a full traversal, and
// with a static split the warps that drew cheap rays left a quarter of k_shadow's warp slots idle (ncu: achieved
// occupancy 29 % of a theoretical 37.5 %).
template <bool LOOP, class Tracer>
__device__ __forceinline__ void phase_shadow(const SceneView& s, const Tracer& tr, RayCounts& rc, const long long n_hits,
                                             const ShadowRec* shq, const unsigned* shq_count, unsigned char* occl,
                                             float4* att4, unsigned long long* work = nullptr, const bool from_back = false) {
    long long total = 0;
    for (int l = 0; l < s.n_lights; l++) total += shq_count[l];
    long long f0 = 0, fstep = 0;
    const bool dynamic = work != nullptr && total >= (long long)gridDim.x * blockDim.x;
    if (!dynamic && !sparse_lanes(total, f0, fstep)) return;
    const int lane = threadIdx.x & 31;
    for (long long f = f0;; f += fstep) {
        if (dynamic) {
            unsigned long long b = 0;
            if (lane == 0) b = atomicAdd(work, 32ull);
            b = __shfl_sync(0xffffffffu, b, 0);
            if ((long long)b >= total) break;
            if (from_back) b = (unsigned long long)(((total + 31) & ~31ll) - 32) - b;  // the queues' last entries first
            f = (long long)b + lane;
            if (f >= total) continue;
        } else if (f >= total) {
            break;
        }
        int l = 0;
        long long k = f;
        for (unsigned c; k >= (long long)(c = shq_count[l]); l++) k -= c;  // light whose queue holds entry f
        ShadowRec q = shq[(size_t)l * n_hits + k];
        if (q.o_maxt.w < 0.f) continue;  // wave-0 hole: culled light
        size_t j = (size_t)l * n_hits + (size_t)__float_as_int(q.d.w);
        rtx::Ray r;
        r.o = rtx::mk3(q.o_maxt.x, q.o_maxt.y, q.o_maxt.z);
        r.d = rtx::mk3(q.d.x, q.d.y, q.d.z);
        if (LOOP) {
            att4[j] = light_attenuate_loop(s, tr, s.lights[l], r, q.o_maxt.w, rc);
        } else {
            rc.shadow++;
            occl[j] = tr.occluded(r, q.o_maxt.w) ? 1 : 0;
        }
    }
}

template <int MODE, bool STATS, bool LOOP>
__global__ void __launch_bounds__(kThreads, RT_SHADOW_MIN_BLOCKS) k_shadow(SceneView s, unsigned long long* ctr, int ctr_n, const ShadowRec* shq,
                                                       const unsigned* shq_count, unsigned char* occl, float4* att4,
                                                       unsigned long long* work, int from_back) {
    TravStats ts = {0, 0, 0};
    RayCounts rc = {0, 0, 0};
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, STATS ? &ts : nullptr);
    phase_shadow<LOOP>(s, tr, rc, (long long)(unsigned)ctr[ctr_n], shq, shq_count, occl, att4, work, from_back != 0);
    flush_counts(ctr, rc, ts, STATS);
}

// Wave 0, opaque scenes: the shadow-ray queue of a light is slot-indexed like the tile-aligned hit queue, so 32 consecutive
// entries are the shadow rays of ONE 8x4 tile towards ONE light (holes: misses and culled lights). Walked together by the
// warp-coherent traversal of rt_beam.cuh instead of 32 per-lane occlusion queries; same flags out.
template <bool STATS, bool FAST = false>
__global__ void __launch_bounds__(kThreads, RT_SHADOW_MIN_BLOCKS) k_shadow_beam(SceneView s, unsigned long long* ctr, int ctr_n,
                                                                               const ShadowRec* shq, unsigned char* occl,
                                                                               unsigned long long* work, int from_back) {
    TravStats ts = {0, 0, 0};
    BeamStats bs = {0};
    RayCounts rc = {0, 0, 0};
    const long long n_hits = (long long)(unsigned)ctr[ctr_n];  // queue slots: a multiple of 32
    const long long total = n_hits * s.n_lights;
    const int lane = threadIdx.x & 31;
    for (;;) {
        unsigned long long b = 0;
        if (lane == 0) b = atomicAdd(work, 32ull);
        b = __shfl_sync(0xffffffffu, b, 0);
        if ((long long)b >= total) break;
        if (from_back) b = (unsigned long long)(total - 32) - b;  // the queue's last groups (bottom of the image) first
        const long long f = (long long)b + lane;
        const long long l = (long long)b / n_hits;  // warp-uniform: n_hits is a multiple of 32
        const ShadowRec q = shq[f];
        const bool trace = q.o_maxt.w >= 0.f;       // not a hole
        rtx::Ray r;
        r.o = rtx::mk3(q.o_maxt.x, q.o_maxt.y, q.o_maxt.z);
        r.d = rtx::mk3(q.d.x, q.d.y, q.d.z);
        Hit h;
        unsigned long long t0 = 0;
        const TravStats ts0 = ts;
        const unsigned nf0 = bs.node_fetches;
        if (STATS) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
        const bool blocked = beam_query<STATS, FAST>(s, r, trace, true, true, q.o_maxt.w, h, ts, bs);
        if (STATS) {  // longest single work unit of the launch (RtFrameStats.ms_wave_phase[3] for frames without bounce waves)
            unsigned long long t1;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
            // diagnostics of that unit (racy, last writer wins): [0] = queue position, [1] = node fetches * 10000 + most exact
            // tests of a lane, [2] = most triangle filters of a lane (overwrites k_tile<1>'s longest tile)
            unsigned pf = ts.prims - ts0.prims, ex = ts.exact - ts0.exact;
            for (int o = 16; o; o >>= 1) { pf = max(pf, __shfl_xor_sync(0xffffffffu, pf, o)); ex = max(ex, __shfl_xor_sync(0xffffffffu, ex, o)); }
            if (lane == 0 && atomicMax(&ctr[C_PHASE_NS + 3], t1 - t0) < t1 - t0) {
                ctr[C_PHASE_NS + 0] = b;
                ctr[C_PHASE_NS + 1] = (unsigned long long)(bs.node_fetches - nf0) * 10000ull + ex;  // small integers: exact as floats
                ctr[C_PHASE_NS + 2] = pf;
            }
        }
        if (trace) {
            rc.shadow++;
            occl[(size_t)l * n_hits + (size_t)__float_as_int(q.d.w)] = blocked ? 1 : 0;
        }
    }
    flush_counts(ctr, rc, ts, STATS);
    if (STATS) {
        unsigned f = bs.node_fetches;
        for (int o = 16; o; o >>= 1) f += __shfl_xor_sync(0xffffffffu, f, o);
        if (lane == 0 && f) atomicAdd(&ctr[C_FETCHES], (unsigned long long)f);
    }
}

// What Light::attenuate returned for (light l, hit i of n), from whichever k_shadow variant ran.
struct AttenLookup {
    const unsigned char* occl;
    const float4* att4;
    size_t n, i;
    __device__ __forceinline__ float4 operator()(int l, const LightRec& L) const {
        size_t j = (size_t)l * n + i;
        if (att4) return att4[j];
        return occl[j] ? rtx::mk4(0.f, 0.f, 0.f, 0.f) : L.color;
    }
};

// SECONDARY == true: hits on reflective / refractive materials become paths (their first
// RayFrame transitions run here) and are queued for the bounce waves.
template <bool SECONDARY>
__global__ void __launch_bounds__(kThreads, RT_RESOLVE_MIN_BLOCKS) k_resolve(SceneView s, unsigned long long* ctr, const HitRec* hitq,
                                                     const unsigned char* occl, const float4* att4, uint32_t* canvas,
                                                     unsigned char* paths, int path_stride, int* queue) {
    const int n = (int)ctr[C_HITQ];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        HitRec rec = hitq[i];
        if (rec.inst < 0) continue;  // a missed pixel inside a tile-aligned group (k_primary wrote its black)
        Hit h;
        h.t = rec.t; h.t_local = rec.t_local; h.u = rec.u; h.v = rec.v; h.inst = rec.inst; h.tri = rec.tri;
        rtx::Ray r = rtx::camera_ray(s.cam, (float)(rec.xy & 0xffff), (float)((unsigned)rec.xy >> 16));
        int mat_id = s.tris[h.tri].mat;
        const MatRec& m = s.mats[mat_id];
        AttenLookup att = {occl, att4, (size_t)n, (size_t)i};
        if (SECONDARY && (m.reflective || m.refractive)) {
            int slot = warp_alloc(&ctr[C_PATHS]);
            PathHead* head;
            Frame* frames = path_frames(paths, path_stride, slot, &head);
            Frame top;  // the root frame starts in registers; path_advance_core stores what stays on the stack
            frame_init_root(top, r, s.depth);
            int stack_top = 0, isect_mat = mat_id;
            float4 acc = rtx::mk4(0.f, 0.f, 0.f, 0.f);
            bool alive = path_advance_core(
                s, [&](const rtx::Ray& ray, float t, float3 nn, const MatRec& mm, float4 kd) { return illuminate_resolved(s, ray, t, nn, mm, kd, att); },
                frames, stack_top, isect_mat, acc, true, h, top, true);
            if (alive) {
                PathHead hd;
                hd.acc = acc; hd.out_index = rec.out_index; hd.stack_top = stack_top; hd.isect_mat = isect_mat;
                hd.pending = 0; hd.hit = h; hd.pad0 = hd.pad1 = 0;
                *head = hd;
                queue[warp_alloc(&ctr[C_QUEUE_A])] = slot;
            } else {
                canvas[rec.out_index] = rtx::pack_color(acc);
            }
            continue;
        }
        // no secondary rays: acc = 0 + 1 * illuminate, clamp, pack (src/raytracer.cu:36-41)
        float3 nrm = hit_normal(s, h, r);
        float4 c = illuminate_resolved(s, r, h.t, nrm, m, diffuse_colour(s, m, h.tri, h.u, h.v), att);
        canvas[rec.out_index] = rtx::pack_color(c);
    }
}

// waves >= 1: closest hit of every live path's pending ray (frames[stack_top].ray)
template <class Tracer>
__device__ __forceinline__ void phase_path_trace(const SceneView& s, const Tracer& tr, RayCounts& rc, const int n,
                                                 const int* queue_in, unsigned char* paths, int path_stride, HitRec* whit) {
    long long i0, istep;
    if (!sparse_lanes(n, i0, istep)) return;
    for (int i = (int)i0; i < n; i += (int)istep) {
        int slot = queue_in[i];
        PathHead* head;
        Frame* frames = path_frames(paths, path_stride, slot, &head);
        rtx::Ray r = frames[head->stack_top].ray;
        Hit h;
        rc.secondary++;
        bool hit = tr.closest(r, h);
        HitRec rec;
        rec.out_index = slot; rec.xy = 0; rec.inst = hit ? h.inst : -1; rec.tri = h.tri;
        rec.t = h.t; rec.u = h.u; rec.v = h.v; rec.t_local = h.t_local;
        whit[i] = rec;
    }
}

// first step of a wave: reset what the later steps of this wave append to (their previous readers,
// the steps of the wave before, have completed)
__device__ __forceinline__ void wave_reset(const SceneView& s, unsigned long long* ctr, int ctr_out, unsigned* shq_count) {
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) ctr[ctr_out] = 0ull;
        for (int l = threadIdx.x; l < s.n_lights; l += blockDim.x) shq_count[l] = 0u;
    }
}

// waves >= 1: shade the hit from the traced shadow rays, run the RayFrame transitions, requeue or finish
__device__ __forceinline__ void phase_path_advance(const SceneView& s, const int n, unsigned long long* live_out,
                                                   const HitRec* whit, const unsigned char* occl, const float4* att4,
                                                   int* queue_out, uint32_t* canvas, unsigned char* paths, int path_stride) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        HitRec rec = whit[i];
        int slot = rec.out_index;
        PathHead* head;
        Frame* frames = path_frames(paths, path_stride, slot, &head);
        float4 acc = head->acc;
        int stack_top = head->stack_top, isect_mat = head->isect_mat;
        Hit h;
        h.t = rec.t; h.t_local = rec.t_local; h.u = rec.u; h.v = rec.v; h.inst = rec.inst; h.tri = rec.tri;
        AttenLookup att = {occl, att4, (size_t)n, (size_t)i};
        bool alive = path_advance_with(
            s, [&](const rtx::Ray& ray, float t, float3 nn, const MatRec& mm, float4 kd) { return illuminate_resolved(s, ray, t, nn, mm, kd, att); },
            frames, stack_top, isect_mat, acc, rec.inst >= 0, h);
        if (alive) {
            head->acc = acc;
            head->stack_top = stack_top;
            head->isect_mat = isect_mat;
            queue_out[warp_alloc(live_out)] = slot;
        } else {
            canvas[head->out_index] = rtx::pack_color(acc);
        }
    }
}

// All bounce waves in one cooperative launch (every block resident): the four steps of a wave run
// as phases separated by grid barriers, and the loop ends on the device when a wave leaves no
// live path. At 640x480 the per-wave launches were host-bound (16 launches, ~20 us each, around
// kernels that run for 3-10 us); this removes them and the per-wave read-back.
__device__ __forceinline__ void wave_barrier(unsigned* counter, unsigned n_blocks, unsigned& epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        epoch++;
        __threadfence();
        atomicAdd(counter, 1u);
        const unsigned target = epoch * n_blocks;
        while (*(volatile unsigned*)counter < target) {}
        __threadfence();
    }
    __syncthreads();
}

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

struct WaveArgs {
    unsigned long long* ctr;
    int* queue_a;
    int* queue_b;
    HitRec* whit;
    ShadowRec* shq;
    unsigned* shq_count;
    unsigned char* occl;
    float4* att4;
    uint32_t* canvas;
    unsigned char* paths;
    int path_stride;
    int max_waves;
    unsigned* barrier;  // zeroed before the launch
};

template <int MODE, bool STATS, bool LOOP>
__global__ void __launch_bounds__(kThreads, RT_WAVES_MIN_BLOCKS) k_bounce_waves(SceneView s, WaveArgs a) {
    TravStats ts = {0, 0, 0};
    RayCounts rc = {0, 0, 0};
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, STATS ? &ts : nullptr);
    unsigned epoch = 0, culled = 0;
    int in = C_QUEUE_A, out = C_QUEUE_B;
    int* qin = a.queue_a;
    int* qout = a.queue_b;
    int waves = 0;
    for (;;) {
        // written by the previous wave's advance phase (or k_resolve); every thread sees the same value
        const int n = (int)*(volatile unsigned long long*)&a.ctr[in];
        if (n == 0 || waves >= a.max_waves) break;
        // STATS: block 0 accumulates the time between barrier completions per phase (ns)
        const bool clock = STATS && blockIdx.x == 0 && threadIdx.x == 0;
        unsigned long long t0 = 0, t1;
        if (clock) t0 = global_ns();
        wave_reset(s, a.ctr, out, a.shq_count);
        phase_path_trace(s, tr, rc, n, qin, a.paths, a.path_stride, a.whit);
        wave_barrier(a.barrier, gridDim.x, epoch);
        if (clock) { t1 = global_ns(); a.ctr[C_PHASE_NS + 0] += t1 - t0; t0 = t1; }
        phase_shadow_gen<true>(s, n, a.whit, a.shq, a.shq_count, a.occl, a.att4, a.paths, a.path_stride, culled);
        wave_barrier(a.barrier, gridDim.x, epoch);
        if (clock) { t1 = global_ns(); a.ctr[C_PHASE_NS + 1] += t1 - t0; t0 = t1; }
        phase_shadow<LOOP>(s, tr, rc, (long long)n, a.shq, a.shq_count, a.occl, a.att4);
        wave_barrier(a.barrier, gridDim.x, epoch);
        if (clock) { t1 = global_ns(); a.ctr[C_PHASE_NS + 2] += t1 - t0; t0 = t1; }
        phase_path_advance(s, n, &a.ctr[out], a.whit, a.occl, a.att4, qout, a.canvas, a.paths, a.path_stride);
        wave_barrier(a.barrier, gridDim.x, epoch);
        if (clock) { t1 = global_ns(); a.ctr[C_PHASE_NS + 3] += t1 - t0; }
        int t = in; in = out; out = t;
        int* q = qin; qin = qout; qout = q;
        waves++;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) a.ctr[C_WAVES] = (unsigned long long)waves;
    flush_counts(a.ctr, rc, ts, STATS);
    flush_culled(s, a.ctr, culled);
}

// ---------------------------------------------------------------- wave 0 in ONE kernel: k_tile
// Camera rays, closest hit, shadow rays, Phong and the pixel store for one 8x4 tile per warp, with the warp-coherent
// traversal of rt_beam.cuh for the camera rays and for the shadow rays of each light. Replaces, for scenes whose shadow
// rays are plain occlusion queries (no material refracts), the chain k_primary -> hit queue -> k_shadow_gen -> shadow
// queues -> k_shadow -> occlusion flags -> k_resolve: the 32-byte hit records, the 32-byte shadow records and the flags
// (0.7 GB of the 1.07 GB of HBM traffic of a 4K world16 frame) are never written, three kernel launches and their tails
// disappear, and what reaches HBM is 4 bytes per pixel. The warp keeps the tile from ray generation to the pixel: the
// tile-aligned queues of the split pipeline already gave every warp of the shading kernels the hits of ONE tile, so the
// lanes idle here (missed pixels of a partly covered tile) idled there as well.
// Arithmetic: the same rt_shade.cuh / rt_scene.cuh functions in the same order as k_shadow_gen + k_resolve, so frames are
// bit-identical to the split pipeline's. SECONDARY: hits on reflective materials start a path (RayFrame stack) exactly as
// k_resolve<true> does and the bounce waves continue from there.
// PEERS > 0: multi-GPU — the pixel goes straight into the row-major canvases of `PEERS` ranks over NVLink peer
// mappings (the gathering rank's, or every rank's) instead of a local slab that a second kernel pushes; the last block
// to finish signals the frame's epoch to their mailboxes (rt_render with RtRenderParams.peers).
struct PeerCanvases { uint32_t* p[RT_MAX_PEERS]; int n; };
struct PushSignal { long long mailbox_offset; unsigned epoch; unsigned* done_blocks; const unsigned* epoch_ptr; };  // epoch_ptr != null: the epoch is read from there (replayed graphs)
struct PeerOut { PeerCanvases pc; PushSignal sg; int rank; };

// PHASE 0: the whole tile in one kernel. PHASE 1 + PHASE 2: the same code cut at the hit — k_tile<1> finds the closest
// hits and appends them tile-aligned (32 slots per tile with a hit, lane l -> slot l) to the hit queue, k_tile<2> pulls
// those groups and shades them. Two launches and 64 B of queue traffic per hit slot, but each kernel's hot loop fits
// the instruction cache and runs with fewer registers (measured: profiles/r2_summary.md).
#ifndef RT_TILE_MIN_BLOCKS
#define RT_TILE_MIN_BLOCKS 2
#endif
#ifndef RT_TILE1_MIN_BLOCKS
#define RT_TILE1_MIN_BLOCKS 4   /* 64 registers, 60 B of spills: k_tile<1> 0.727 -> 0.706 ms against 3 blocks / 80 registers (profiles/r2_summary.md) */
#endif
#ifndef RT_TILE2_MIN_BLOCKS
#define RT_TILE2_MIN_BLOCKS 2   /* 128 registers: at 3 blocks (80) the shading phase spills 324 B and runs 6 % slower */
#endif
template <bool STATS, bool SECONDARY, bool PEERS, int PHASE, bool FAST = false>
__global__ void __launch_bounds__(kThreads, PHASE == 0 ? RT_TILE_MIN_BLOCKS : PHASE == 1 ? RT_TILE1_MIN_BLOCKS : RT_TILE2_MIN_BLOCKS)
k_tile(SceneView s, TileMap tm, unsigned long long* ctr, uint32_t* canvas, PeerOut po, unsigned char* paths, int path_stride, int* queue,
       HitRec* hitq) {
    TravStats ts = {0, 0, 0};
    BeamStats bs = {0};
    RayCounts rc = {0, 0, 0};
    unsigned culled = 0, n_hit = 0;
    const int lane = threadIdx.x & 31;
    const int n_work = tm.n_work();
    auto store = [&](int out, uint32_t v) {
        if (PEERS) {
            for (int k = 0; k < po.pc.n; k++) po.pc.p[k][out] = v;
        } else {
            canvas[out] = v;
        }
    };
    const unsigned n_groups = PHASE == 2 ? (unsigned)ctr[C_HITQ] >> 5 : 0u;  // 32-slot hit groups left by k_tile<1>
    for (;;) {
        int q = 0;
        if (lane == 0) q = (int)atomicAdd(&ctr[PHASE == 2 ? C_WORK1 : C_WORK0], 1ull);
        q = __shfl_sync(0xffffffffu, q, 0);
        int x, y, out;
        rtx::Ray r;
        r.o = s.cam.pos; r.d = s.cam.f;
        Hit h;
        bool hit;
        if (PHASE != 2) {
            if (q >= n_work) break;
            // Bottom of the image first: tiles are enumerated from the top (sky: one node test) to the bottom (terrain near the
            // camera: the most expensive tiles). A persistent kernel that meets its heaviest units last ends with a few warps
            // still walking them (SMs finished up to 0.04 ms apart; that imbalance does not shrink when the frame is sharded
            // over more GPUs). Reversed, the cheap units come last and even the end out; the hit queue — filled in completion
            // order — then hands the shadow kernel its heavy units first as well. RT_TILE_ORDER=0 restores the old order.
            if (tm.reverse == 1) q = n_work - 1 - q;
            else if (tm.reverse == 2) { q += n_work >> 1; if (q >= n_work) q -= n_work; }  // lower half of the image first, in natural order
            const bool valid = tm.locate(q, lane, x, y, out);
            if (valid) {
                r = rtx::camera_ray(s.cam, (float)x, (float)y);
                rc.primary++;
            }
            unsigned long long t0 = 0;
            if (STATS) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
            hit = beam_query<STATS, FAST>(s, r, valid, false, true, INFINITY, h, ts, bs);
            if (STATS) {  // longest single tile (ms_wave_phase[2])
                unsigned long long t1;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
                if (lane == 0) atomicMax(&ctr[C_PHASE_NS + 2], t1 - t0);
            }
            if (valid && !hit) store(out, 0u);  // pack_color(0,0,0,0)
            const unsigned hm = __ballot_sync(0xffffffffu, hit);
            if (!hm) continue;
            if (PHASE == 1) {
                // tile-aligned append (k_primary's): ctr[C_HITQ] low word = queue slots, high word = hits
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(&ctr[C_HITQ], 32ull | ((unsigned long long)__popc(hm) << 32));
                base = __shfl_sync(0xffffffffu, base, 0) & 0xffffffffull;
                HitRec rec;
                rec.out_index = out; rec.xy = x | (y << 16); rec.inst = hit ? h.inst : -1; rec.tri = h.tri;
                rec.t = h.t; rec.u = h.u; rec.v = h.v; rec.t_local = h.t_local;
                hitq[base + (unsigned)lane] = rec;
                continue;
            }
            if (hit) n_hit++;
        } else {
            if (PHASE == 2 && (unsigned)q >= n_groups) break;
            const HitRec rec = hitq[(size_t)q * 32 + lane];
            hit = rec.inst >= 0;
            out = rec.out_index; x = rec.xy & 0xffff; y = (int)((unsigned)rec.xy >> 16);
            h.t = rec.t; h.t_local = rec.t_local; h.u = rec.u; h.v = rec.v; h.inst = rec.inst; h.tri = rec.tri;
            if (hit) r = rtx::camera_ray(s.cam, (float)x, (float)y);
        }
        // ---- shading of the tile's hits: k_shadow_gen + k_shadow + k_resolve in place. illuminate() (phong.cu:35-53):
        // Ke + Ka * ambience, then one Phong term per light in light order — accumulated as the lights' shadow rays are
        // answered, in the order and with the operations of illuminate_resolved().
        float3 nrm = rtx::mk3(0.f, 0.f, 0.f), hit_pos = rtx::mk3(0.f, 0.f, 0.f);
        int mat_id = 0;
        if (hit) {
            mat_id = s.tris[h.tri].mat;
            nrm = hit_normal(s, h, r);
            hit_pos = rtx::ray_at(r, h.t);
        }
        const MatRec& m = s.mats[mat_id];
        float4 kd = rtx::mk4(0.f, 0.f, 0.f, 0.f), sum = kd;
        if (hit) {
            kd = diffuse_colour(s, m, h.tri, h.u, h.v);
            sum = rtx::add4(m.Ke, rtx::mul4(m.Ka, s.ambience));
        }
        for (int l = 0; l < s.n_lights; l++) {
            bool trace = false;
            rtx::Ray sr;
            sr.o = hit_pos; sr.d = r.d;
            float3 dir = r.d;
            float max_t = 0.0f, da = 1.0f;
            if (hit) {
                light_setup(s, s.lights[l], hit_pos, dir, max_t, da);
                if (s.cull_unlit && light_unlit(m, nrm, dir)) {
                    culled++;  // zero Phong weight: the term is +0 whatever the shadow ray says (rt_shade.cuh light_unlit)
                } else {
                    trace = true;
                    sr = shadow_ray(hit_pos, dir);
                    rc.shadow++;
                }
            }
            Hit sh;
            const bool blocked = beam_query<STATS, FAST>(s, sr, trace, true, true, max_t, sh, ts, bs);
            if (trace) {
                const LightRec& L = s.lights[l];
                const float4 a = blocked ? rtx::mk4(0.f, 0.f, 0.f, 0.f) : L.color;
                const float4 incoming = L.kind == 0 ? rtx::scale4(da, a) : a;
                sum = rtx::add4(sum, phong(m, kd, nrm, incoming, r.d, dir));
            }
        }
        if (!hit) continue;
        if (SECONDARY && (m.reflective || m.refractive)) {
            // the pixel continues: first RayFrame transitions here (k_resolve<true>), then the bounce waves
            int slot = warp_alloc(&ctr[C_PATHS]);
            PathHead* head;
            Frame* frames = path_frames(paths, path_stride, slot, &head);
            Frame top;
            frame_init_root(top, r, s.depth);
            int stack_top = 0, isect_mat = mat_id;
            float4 acc = rtx::mk4(0.f, 0.f, 0.f, 0.f);
            bool alive = path_advance_core(
                s, [&](const rtx::Ray&, float, float3, const MatRec&, float4) { return sum; },  // illuminate() of this hit
                frames, stack_top, isect_mat, acc, true, h, top, true);
            if (alive) {
                PathHead hd;
                hd.acc = acc; hd.out_index = out; hd.stack_top = stack_top; hd.isect_mat = isect_mat;
                hd.pending = 0; hd.hit = h; hd.pad0 = hd.pad1 = 0;
                *head = hd;
                queue[warp_alloc(&ctr[C_QUEUE_A])] = slot;
            } else {
                store(out, rtx::pack_color(acc));
            }
            continue;
        }
        // no secondary rays: acc = 0 + 1 * illuminate, clamp, pack (src/raytracer.cu:36-41)
        store(out, rtx::pack_color(sum));
    }
    flush_counts(ctr, rc, ts, STATS);
    flush_culled(s, ctr, culled);
    {
        unsigned v = n_hit, f = bs.node_fetches;
        for (int o = 16; o; o >>= 1) { v += __shfl_xor_sync(0xffffffffu, v, o); f += __shfl_xor_sync(0xffffffffu, f, o); }
        if (PHASE == 0 && lane == 0 && v) atomicAdd(&ctr[C_HITQ], (unsigned long long)v << 32);  // high word = hits (n_hits)
        if (STATS && lane == 0 && f) atomicAdd(&ctr[C_FETCHES], (unsigned long long)f);
    }
    if (PEERS && PHASE != 1) {
        // every block: make its peer stores visible system-wide, then count itself; the last block signals
        __shared__ bool s_last;
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned prev = atomicAdd(po.sg.done_blocks, 1u);
            s_last = prev + 1u == gridDim.x;
        }
        __syncthreads();
        if (s_last) {
            __threadfence_system();
            if ((int)threadIdx.x < po.pc.n) {
                volatile unsigned* box = (volatile unsigned*)(po.pc.p[threadIdx.x] + po.sg.mailbox_offset);
                box[po.rank] = po.sg.epoch_ptr ? *(volatile const unsigned*)po.sg.epoch_ptr : po.sg.epoch;
            }
            if (threadIdx.x == 0) *po.sg.done_blocks = 0u;  // ready for the next frame
        }
    }
}

// ---------------------------------------------------------------- waves >= 1 with ray sorting (RT_RENDER_SORT_RAYS)
// One launch per phase and wave, and between the waves the live queue is sorted by a coherence key: 21-bit Morton code
// of the pending ray's origin inside the scene bounds (7 bits per axis), then the direction octant (24 bits, three radix
// passes). Neighbouring lanes then trace rays that start in the same cell and head the same way. The order of the queue
// does not enter any result (whit / occl are indexed by queue position within a wave).
__device__ __forceinline__ unsigned spread7(unsigned v) {  // 7 bits -> every third bit
    v &= 0x7fu;
    v = (v | (v << 8)) & 0x0000700fu;
    v = (v | (v << 4)) & 0x000430c3u;
    v = (v | (v << 2)) & 0x00049249u;
    return v;
}
__global__ void k_path_keys(const int* __restrict__ queue, int n, const unsigned char* paths, int path_stride,
                            const float* __restrict__ bounds6, unsigned long long* __restrict__ keys, int* __restrict__ vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int slot = queue[i];
    PathHead* head;
    Frame* frames = path_frames((unsigned char*)paths, path_stride, slot, &head);
    rtx::Ray r = frames[head->stack_top].ray;
    float o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z};
    unsigned q[3], oct = 0;
    for (int k = 0; k < 3; k++) {
        float ext = bounds6[3 + k] - bounds6[k];
        float t = ext > 0.0f ? (o[k] - bounds6[k]) / ext : 0.0f;
        q[k] = (unsigned)fminf(fmaxf(t * 128.0f, 0.0f), 127.0f);
        oct |= (d[k] < 0.0f ? 1u : 0u) << k;
    }
    unsigned morton = (spread7(q[0]) << 2) | (spread7(q[1]) << 1) | spread7(q[2]);
    keys[i] = ((unsigned long long)morton << 3) | oct;
    vals[i] = slot;
}
template <int MODE, bool STATS>
__global__ void __launch_bounds__(kThreads, RT_WAVES_MIN_BLOCKS) k_wave_trace(SceneView s, unsigned long long* ctr, int ctr_in, const int* queue_in,
                                                         unsigned char* paths, int path_stride, HitRec* whit) {
    TravStats ts = {0, 0, 0};
    RayCounts rc = {0, 0, 0};
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, STATS ? &ts : nullptr);
    phase_path_trace(s, tr, rc, (int)ctr[ctr_in], queue_in, paths, path_stride, whit);
    flush_counts(ctr, rc, ts, STATS);
}
__global__ void __launch_bounds__(kThreads, RT_WAVES_MIN_BLOCKS) k_wave_advance(SceneView s, unsigned long long* ctr, int ctr_in, int ctr_out,
                                                           const HitRec* whit, const unsigned char* occl, const float4* att4,
                                                           int* queue_out, uint32_t* canvas, unsigned char* paths, int path_stride) {
    phase_path_advance(s, (int)ctr[ctr_in], &ctr[ctr_out], whit, occl, att4, queue_out, canvas, paths, path_stride);
}

// ---------------------------------------------------------------- waves >= 1, fused variant (RT_RENDER_FUSED_SHADE)
template <int MODE, bool STATS>
__global__ void __launch_bounds__(kThreads) k_bounce(SceneView s, unsigned long long* ctr, int ctr_in, int ctr_out,
                                                    const int* queue_in, int* queue_out, uint32_t* canvas,
                                                    unsigned char* paths, int path_stride) {
    TravStats ts = {0, 0, 0};
    RayCounts rc = {0, 0, 0};
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, STATS ? &ts : nullptr);
    const int n = (int)ctr[ctr_in];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int slot = queue_in[i];
        PathHead* head = (PathHead*)(paths + (size_t)slot * path_stride);
        Frame* frames = (Frame*)(head + 1);
        float4 acc = head->acc;
        int stack_top = head->stack_top, isect_mat = head->isect_mat;
        Hit h;
        bool hit;
        if (head->pending) {
            h = head->hit;
            hit = true;
            head->pending = 0;
        } else {
            rtx::Ray r = frames[stack_top].ray;
            rc.secondary++;
            hit = tr.closest(r, h);
        }
        bool alive = path_advance(s, tr, frames, stack_top, isect_mat, acc, hit, h, rc);
        if (alive) {
            head->acc = acc;
            head->stack_top = stack_top;
            head->isect_mat = isect_mat;
            int qi = warp_alloc(&ctr[ctr_out]);
            queue_out[qi] = slot;
        } else {
            canvas[head->out_index] = rtx::pack_color(acc);
        }
    }
    flush_counts(ctr, rc, ts, STATS);
}

// ---------------------------------------------------------------- one thread per pixel, whole ray tree
// Used for `-r` (RT_RENDER_NO_BVH), for RT_SEMANTICS_CPU and by rt_debug_cast.
template <int MODE, int SEM>
__global__ void __launch_bounds__(kThreads) k_mega(SceneView s, TileMap tm, unsigned long long* ctr, uint32_t* canvas) {
    TravStats ts = {0, 0, 0};
    RayCounts rc = {0, 0, 0};
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, nullptr);
    const int lane = threadIdx.x & 31;
    const int n_work = tm.n_work();
    for (;;) {
        int q = 0;
        if (lane == 0) q = (int)atomicAdd(&ctr[C_WORK0], 1ull);
        q = __shfl_sync(0xffffffffu, q, 0);
        if (q >= n_work) break;
        int x, y, out;
        if (!tm.locate(q, lane, x, y, out)) continue;
        rtx::Ray r = rtx::camera_ray(s.cam, (float)x, (float)y);
        float4 c;
        if (SEM == RT_SEMANTICS_CPU) {
            CpuIsect is;
            is.time = INFINITY; is.norm = rtx::mk3(0.f, 0.f, 0.f); is.mat = 0;
            float last_time = 0.f;
            c = propagate_cpu_rec(s, tr, r, is, last_time, 0, s.depth, 0, rc);
        } else {
            c = propagate_gpu(s, tr, r, rc);
        }
        canvas[out] = rtx::pack_color(c);
    }
    flush_counts(ctr, rc, ts, false);
}

template <int MODE>
__global__ void k_cast_rays(SceneView s, const float* origins, const float* dirs, int n, int* inst, int* tri, float* t) {
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, nullptr);
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    rtx::Ray r = rtx::ray_make(rtx::mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                               rtx::mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
    Hit h;
    bool hit = tr.closest(r, h);
    inst[i] = hit ? h.inst : -1;
    tri[i] = hit ? h.tri - s.meshes[s.insts[h.inst].mesh].tri_begin : -1;
    t[i] = h.t;
}

template <int MODE, int SEM>
__global__ void k_debug(SceneView s, int x, int y, RtHit* out) {
    typename TracerOf<MODE>::T tr = TracerOf<MODE>::make(&s, nullptr);
    RayCounts rc = {0, 0, 0};
    rtx::Ray r = rtx::camera_ray(s.cam, (float)x, (float)y);
    Hit h;
    bool hit = tr.closest(r, h);
    RtHit o;
    o.instance = hit ? h.inst : -1;
    o.triangle = hit ? h.tri - s.meshes[s.insts[h.inst].mesh].tri_begin : -1;
    o.t = h.t;
    float3 n = hit ? hit_normal(s, h, r) : rtx::mk3(0.f, 0.f, 0.f);
    o.normal[0] = n.x; o.normal[1] = n.y; o.normal[2] = n.z;
    float4 c;
    if (SEM == RT_SEMANTICS_CPU) {
        CpuIsect is;
        is.time = INFINITY; is.norm = rtx::mk3(0.f, 0.f, 0.f); is.mat = 0;
        float last_time = 0.f;
        c = propagate_cpu_rec(s, tr, r, is, last_time, 0, s.depth, 0, rc);
    } else {
        c = propagate_gpu(s, tr, r, rc);
    }
    o.color[0] = c.x; o.color[1] = c.y; o.color[2] = c.z; o.color[3] = c.w;
    *out = o;
}

// De-interleave `world` gathered slabs into a row-major canvas (after the NCCL all-gather).
__global__ void k_untile(TileMap tm, long long slab, const uint32_t* gathered, uint32_t* canvas) {
    long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= (long long)tm.W * tm.H) return;
    int x = (int)(p % tm.W), y = (int)(p / tm.W);
    int tx = x / tm.tile_w, ty = y / tm.tile_h;
    int T = ty * tm.tiles_x + tx;
    int rank = T % tm.world, k = T / tm.world;
    int lx = x - tx * tm.tile_w, ly = y - ty * tm.tile_h;
    canvas[p] = gathered[(long long)rank * slab + ((long long)k * tm.tile_h + ly) * tm.tile_w + lx];
}

// Multi-GPU frame exchange without a collective library: every rank stores the pixels of the tiles
// it owns straight into the row-major canvas of every peer (its own included) through NVLink peer
// mappings. One thread moves 4 pixels (16 bytes); a tile row is a contiguous run in the slab and
// in the canvas, so a warp writes whole 128-byte segments. Replaces all-gather + k_untile.
// Exchange barrier over peer memory (no collective library on the frame's data path). Behind the pixels of every
// peer canvas sits a mailbox of RT_MAX_PEERS words; after its last tile store a rank writes the frame's epoch into
// slot [rank] of every rank's mailbox, and k_wait_mailbox on the receiving side spins until all slots carry the epoch.

__device__ __forceinline__ void push_signal_tail(const TileMap& tm, const PeerCanvases& pc, const PushSignal& sg) {
    // every block: make its peer stores visible system-wide, then count itself; the last block signals
    __shared__ bool s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned prev = atomicAdd(sg.done_blocks, 1u);
        s_last = prev + 1u == gridDim.x;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence_system();
    if (threadIdx.x < pc.n) {
        volatile unsigned* box = (volatile unsigned*)(pc.p[threadIdx.x] + sg.mailbox_offset);
        box[tm.rank] = sg.epoch_ptr ? *(volatile const unsigned*)sg.epoch_ptr : sg.epoch;
    }
    if (threadIdx.x == 0) *sg.done_blocks = 0u;  // ready for the next frame
}

// ack.n > 0 (gather to one rank): once every rank has signalled, the gathering rank acknowledges the epoch in each sender's own
// mailbox (slot ack_slot = the gathering rank's id; k_wait_ack is the other end).
__global__ void k_wait_mailbox(const uint32_t* own_canvas, long long mailbox_offset, int n, unsigned epoch, long long timeout_cycles,
                               unsigned* status, const unsigned* epoch_ptr, PeerCanvases ack, int ack_slot) {
    if (epoch_ptr) epoch = *(volatile const unsigned*)epoch_ptr;  // replayed graph: this frame's epoch sits in pinned memory
    volatile const unsigned* box = (volatile const unsigned*)(own_canvas + mailbox_offset);
    // epochs only grow; a peer that died would hang the GPU, so the wait gives up after timeout_cycles. The frame is then
    // incomplete: the slot that timed out is recorded in *status (pinned host memory, bit r = rank r never signalled) and
    // the host turns it into RT_ERR_CUDA at the next synchronisation point of this scene (rt_frame_finish / the
    // blocking rt_push_tiles_signal).
    if ((int)threadIdx.x < n) {
        const long long t0 = clock64();
        while ((int)(box[threadIdx.x] - epoch) < 0) {
            if (clock64() - t0 > timeout_cycles) { atomicOr(status, 1u << threadIdx.x); break; }
        }
    }
    __threadfence_system();
    __syncwarp();  // every slot has been seen (or given up on) before any acknowledgement leaves
    if ((int)threadIdx.x < ack.n && (int)threadIdx.x != ack_slot)
        *((volatile unsigned*)(ack.p[threadIdx.x] + mailbox_offset) + ack_slot) = epoch;
}

// Sender side of a gather to one rank: before this rank stores frame `epoch` into the gathering rank's canvas of that parity
// (which still holds frame epoch - 2), the gathering rank must have finished gathering frame epoch - 1 — its wait kernel for
// that frame runs behind everything its stream did with frame epoch - 2. Without this a sender that never waits for anyone
// could run any number of frames ahead of the receiver. In step the acknowledgement is a frame old when it is looked at.
__global__ void k_wait_ack(const uint32_t* own_canvas0, long long mailbox_offset, int slot, unsigned epoch, long long timeout_cycles,
                           unsigned* status, const unsigned* epoch_ptr) {
    if (threadIdx.x) return;
    if (epoch_ptr) epoch = *(volatile const unsigned*)epoch_ptr;
    volatile const unsigned* box = (volatile const unsigned*)(own_canvas0 + mailbox_offset) + slot;
    const long long t0 = clock64();
    while ((int)(*box - (epoch - 1u)) < 0) {
        if (clock64() - t0 > timeout_cycles) { atomicOr(status, 1u << slot); break; }
    }
    __threadfence_system();
}

__device__ __forceinline__ void push_tile_quad(const TileMap& tm, const uint32_t* slab, const PeerCanvases& pc, long long q, long long n_quads) {
    const int quads_per_row = tm.tile_w >> 2;
    if (q >= n_quads) return;
    int lq = (int)(q % quads_per_row);
    long long row = q / quads_per_row;
    int ly = (int)(row % tm.tile_h), k = (int)(row / tm.tile_h);
    int T = k * tm.world + tm.rank;
    int tx = T % tm.tiles_x, ty = T / tm.tiles_x;
    int x = tx * tm.tile_w + lq * 4, y = ty * tm.tile_h + ly;
    if (y >= tm.H || x >= tm.W) return;
    const uint32_t* src = slab + ((long long)k * tm.tile_h + ly) * tm.tile_w + lq * 4;
    long long dst = (long long)y * tm.W + x;
    if (x + 4 <= tm.W && (tm.W & 3) == 0) {
        uint4 v = *(const uint4*)src;
        for (int r = 0; r < pc.n; r++) *(uint4*)(pc.p[r] + dst) = v;
    } else {
        for (int i = 0; i < 4 && x + i < tm.W; i++) {
            uint32_t v = src[i];
            for (int r = 0; r < pc.n; r++) pc.p[r][dst + i] = v;
        }
    }
}

template <bool SIGNAL>
__global__ void k_push_tiles(TileMap tm, const uint32_t* slab, PeerCanvases pc, PushSignal sg) {
    const int quads_per_row = tm.tile_w >> 2;
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long n_quads = (long long)tm.n_owned * tm.tile_h * quads_per_row;
    push_tile_quad(tm, slab, pc, q, n_quads);
    if (SIGNAL) push_signal_tail(tm, pc, sg);
}
// RT_RENDER_PEER_STORE when the frame could not store its pixels into the peers itself (secondary rays, refractive
// scenes, -r / serial semantics): the rank rendered its tiles into its own row-major canvas; this kernel forwards them
// to the other target canvases and signals all targets.
__global__ void k_forward_tiles(TileMap tm, const uint32_t* own, PeerCanvases pc, PushSignal sg) {
    const int quads_per_row = tm.tile_w >> 2;
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n_quads = (long long)tm.n_owned * tm.tile_h * quads_per_row;
    if (q < n_quads) {
        const int lq = (int)(q % quads_per_row);
        const long long row = q / quads_per_row;
        const int ly = (int)(row % tm.tile_h), k = (int)(row / tm.tile_h);
        const int T = k * tm.world + tm.rank;
        const int tx = T % tm.tiles_x, ty = T / tm.tiles_x;
        const int x = tx * tm.tile_w + lq * 4, y = ty * tm.tile_h + ly;
        if (y < tm.H && x < tm.W) {
            const long long dst = (long long)y * tm.W + x;
            if (x + 4 <= tm.W && (tm.W & 3) == 0) {
                const uint4 v = *(const uint4*)(own + dst);
                for (int r = 0; r < pc.n; r++) if (pc.p[r] != own) *(uint4*)(pc.p[r] + dst) = v;
            } else {
                for (int i = 0; i < 4 && x + i < tm.W; i++) {
                    const uint32_t v = own[dst + i];
                    for (int r = 0; r < pc.n; r++) if (pc.p[r] != own) pc.p[r][dst + i] = v;
                }
            }
        }
    }
    push_signal_tail(tm, pc, sg);
}

// Resident-working-set read bandwidth (the L2 roofline denominator).
__global__ void k_l2_read(const uint4* buf, long long n16, int iters, unsigned* sink) {
    unsigned acc = 0;
    long long stride = (long long)gridDim.x * blockDim.x;
    for (int it = 0; it < iters; it++)
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
            uint4 v = __ldcg(buf + i);
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345678u) *sink = acc;
}

// Instance records are written straight into a pinned, grow-only staging buffer and copied asynchronously (a pageable
// temporary cost 5 ms per call at 314 k instances). ev[10] guards the staging buffer against the previous copy.
static int upload_instances(RtScene* sc) {
    const size_t n = sc->instances.size();
    const size_t bytes = sizeof(InstRec) * std::max<size_t>(n, 1);
    if (sc->d_insts.reserve(bytes)) return RT_ERR_OUT_OF_MEMORY;
    if (sc->h_inst_cap < bytes) {
        if (sc->h_inst) { RT_CUDA(cudaEventSynchronize(sc->ev[10])); RT_CUDA(cudaFreeHost(sc->h_inst)); sc->h_inst = nullptr; sc->h_inst_cap = 0; }
        RT_CUDA(cudaHostAlloc(&sc->h_inst, bytes + bytes / 2, cudaHostAllocDefault));
        sc->h_inst_cap = bytes + bytes / 2;
    } else {
        RT_CUDA(cudaEventSynchronize(sc->ev[10]));
    }
    rth::fill_instances(sc->instances.data(), (int)n, (InstRec*)sc->h_inst);
    {
        bool ident = sc->meshes_identity;
        const InstRec* ir = (const InstRec*)sc->h_inst;
        for (size_t i = 0; ident && i < n; i++) ident = ir[i].pose.identity_rot != 0;
        sc->all_identity = ident;
    }
    if (n) RT_CUDA(cudaMemcpyAsync(sc->d_insts.p, sc->h_inst, sizeof(InstRec) * n, cudaMemcpyHostToDevice, sc->stream));
    RT_CUDA(cudaEventRecord(sc->ev[10], sc->stream)); /* outside any frame */
    sc->boxes_valid = false;
    sc->bvh.valid = false;
    return RT_OK;
}

template <class T>
static int upload(rth::DevBuf& b, const std::vector<T>& v, cudaStream_t st) {
    if (b.reserve(sizeof(T) * std::max<size_t>(v.size(), 1))) return RT_ERR_OUT_OF_MEMORY;
    if (!v.empty()) RT_CUDA(cudaMemcpyAsync(b.p, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice, st));
    return RT_OK;
}

}  // namespace

namespace rth {

int update_instance_boxes(RtScene* sc) {
    if (sc->boxes_valid) return RT_OK;
    if (sc->n_instances > 0) {
        k_instance_boxes<<<(sc->n_instances + kThreads - 1) / kThreads, kThreads, 0, sc->stream>>>(
            sc->d_insts.as<InstRec>(), sc->n_instances, sc->d_meshes.as<MeshRec>(), sc->d_mesh_box.as<float>());
        sc->kernel_launches++;
        RT_CUDA(cudaGetLastError());
    }
    sc->boxes_valid = true;
    return RT_OK;
}

SceneView make_view(const RtScene* sc, bool with_bvh) {
    SceneView v;
    std::memset(&v, 0, sizeof v);
    v.verts = sc->d_verts.as<float4>();
    v.norms = sc->d_norms.as<float4>();
    v.tris = sc->d_tris.as<TriRec>();
    static const int pre_off = [] { const char* e = std::getenv("RT_TRI_PRE"); return (e && !std::atoi(e)) ? 1 : 0; }();
    v.tri_pre = (sc->d_tri_pre.p && !pre_off) ? sc->d_tri_pre.as<float4>() : nullptr;
    v.meshes = sc->d_meshes.as<MeshRec>();
    v.insts = sc->d_insts.as<InstRec>();
    v.mats = sc->d_mats.as<MatRec>();
    v.lights = sc->d_lights.as<LightRec>();
    if (sc->textured) {
        v.texc = sc->d_texc.as<TexRec>();
        v.atlas = sc->d_atlas.as<float4>();
        v.atlas_w = sc->atlas_w; v.atlas_h = sc->atlas_h;
    }
    v.n_insts = sc->n_instances;
    v.n_lights = sc->n_lights;
    if (with_bvh && sc->bvh.valid) {
        v.n_prims = sc->bvh.n_prims;
        v.nodes = sc->bvh.nodes8.as<uint4>();
        v.prims = sc->bvh.prims8.as<float4>();
    }
    const RtCamera& c = sc->camera;
    v.cam = rtx::camera_basis(c.pose.rot, c.pose.pos, c.global_near, c.unit_to_pixels, c.canvas_width, c.canvas_height);
    v.ambience = make_float4(sc->env.ambience[0], sc->env.ambience[1], sc->env.ambience[2], sc->env.ambience[3]);
    v.dist_c0 = sc->env.dist_atten[0]; v.dist_c1 = sc->env.dist_atten[1]; v.dist_c2 = sc->env.dist_atten[2];
    v.depth = sc->env.depth;
    v.width = sc->width; v.height = sc->height;
    v.any_refractive = sc->any_refractive;
    v.any_secondary = sc->any_secondary;
    // not with refractive materials: a culled ray would have cost the reference an unknown number
    // of Light::attenuate casts, and rays_shadow must stay the reference's count
    v.cull_unlit = (sc->cull_ok && !sc->any_refractive && !sc->trace_unlit && !sc->fused_shade) ? 1 : 0;
    v.one_bits = 0x3f800000u;
    // tolerance mode: the render flag (or RT_FAST_ACCEPT=1 for the parity hooks), only for scenes without rotated instances or
    // meshes — the reference's leaf box of a rotated instance need not contain its geometry (from_local quirk), so there the
    // leaf-box rule can reject hits that lie clearly inside a triangle and the exact path has to decide
    {
        const char* e = std::getenv("RT_FAST_ACCEPT");
        v.fast_accept = ((sc->fast_accept || (e && std::atoi(e))) && sc->all_identity) ? 1 : 0;
    }
    return v;
}

}  // namespace rth

// ================================================================ C ABI
extern "C" {

const char* rt_last_error_string(void) { return rth::g_err.c_str(); }
const char* rt_version(void) { return "rtb200 0.1 (sm_100a)"; }

int rt_device_count(int* out) {
    if (!out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { *out = 0; return rth::cuda_fail(e, "cudaGetDeviceCount", __FILE__, __LINE__); }
    *out = n;
    return RT_OK;
}

int rt_init(int device) {
    int n = 0;
    if (int rc = rt_device_count(&n)) return rc;
    if (n <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_NO_DEVICE; }
    if (device < 0 || device >= n) { set_error("device ordinal out of range"); return RT_ERR_INVALID_ARG; }
    RT_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    RT_CUDA(cudaGetDeviceProperties(&prop, device));
    g_device = device;
    g_sm_count = prop.multiProcessorCount;
    return RT_OK;
}

int rt_scene_create(const RtSceneDesc* d, RtScene** out) {
    if (!d || !out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    *out = nullptr;
    if (g_device < 0) { if (int rc = rt_init(0)) return rc; }
    if (d->n_vertices < 0 || d->n_triangles < 0 || d->n_meshes < 0 || d->n_instances < 0 || d->n_lights < 0 ||
        d->width <= 0 || d->height <= 0 || d->width > 65535 || d->height > 65535) {
        set_error("negative count or canvas size out of range (1..65535)");
        return RT_ERR_INVALID_ARG;
    }
    if ((d->n_vertices && (!d->vertices || !d->normals)) || (d->n_triangles && !d->triangles) || (d->n_meshes && !d->meshes) ||
        (d->n_instances && !d->instances) || (d->n_lights && !d->lights)) {
        set_error("null array with non-zero count");
        return RT_ERR_INVALID_ARG;
    }
    if ((long long)d->width * d->height > (1ll << 30)) {  // pixel indices are 32-bit ints in the kernels (y * W + x)
        set_error("canvas larger than 2^30 pixels");
        return RT_ERR_LIMIT;
    }
    if (d->env.depth < 0 || d->env.depth >= RT_MAX_FRAMES) {  // RayFrame frames[MAX_DEPTH], scene.cu:25,96
        set_error("depth must be in 0..9 (the reference's RayFrame[10] stack)");
        return RT_ERR_LIMIT;
    }
    for (int i = 0; i < d->n_meshes; i++)
        if (d->meshes[i].tri_begin < 0 || d->meshes[i].tri_end < d->meshes[i].tri_begin || d->meshes[i].tri_end > d->n_triangles) { set_error("mesh triangle range out of range"); return RT_ERR_INVALID_ARG; }
    {   // vertex indices of every record a triangle mesh can reach (a sphere's material record carries none)
        std::vector<char> is_sphere_rec((size_t)d->n_triangles, 0);
        for (int i = 0; i < d->n_meshes; i++)
            if (d->meshes[i].kind == RT_HITABLE_SPHERE)
                for (int t = d->meshes[i].tri_begin; t < d->meshes[i].tri_end; t++) is_sphere_rec[t] = 1;
        for (int i = 0; i < d->n_triangles; i++)
            for (int k = 0; k < 3 && !is_sphere_rec[i]; k++)
                if (d->triangles[i].v[k] < 0 || d->triangles[i].v[k] >= d->n_vertices) { set_error("triangle vertex index out of range"); return RT_ERR_INVALID_ARG; }
    }
    for (int i = 0; i < d->n_meshes; i++) {
        const RtMesh& m = d->meshes[i];
        if (m.kind != RT_HITABLE_TRIMESH && m.kind != RT_HITABLE_SPHERE) { set_error("unknown hitable kind"); return RT_ERR_INVALID_ARG; }
        if (m.kind == RT_HITABLE_SPHERE && (m.tri_end - m.tri_begin != 1 || !(m.radius > 0.0f) || !std::isfinite(m.radius))) {
            set_error("a sphere hitable owns exactly one material record and needs a finite radius > 0");
            return RT_ERR_INVALID_ARG;
        }
    }
    for (int i = 0; i < d->n_instances; i++)
        if (d->instances[i].mesh < 0 || d->instances[i].mesh >= d->n_meshes) { set_error("instance mesh index out of range"); return RT_ERR_INVALID_ARG; }

    RtScene* sc = new (std::nothrow) RtScene();
    if (!sc) { set_error("host allocation failed"); return RT_ERR_OUT_OF_MEMORY; }
    auto fail = [&](int rc) { rt_scene_destroy(sc); return rc; };
    sc->sm_count = g_sm_count;
    cudaError_t e = cudaStreamCreateWithFlags(&sc->own_stream, cudaStreamNonBlocking);
    sc->stream = sc->own_stream;
    if (e != cudaSuccess) return fail(rth::cuda_fail(e, "cudaStreamCreate", __FILE__, __LINE__));
    for (int i = 0; i < 12; i++) {
        e = cudaEventCreate(&sc->ev[i]);
        if (e != cudaSuccess) return fail(rth::cuda_fail(e, "cudaEventCreate", __FILE__, __LINE__));
    }
    e = cudaStreamCreateWithFlags(&sc->copy_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) return fail(rth::cuda_fail(e, "cudaStreamCreate", __FILE__, __LINE__));
    for (int c = 0; c < RtScene::kMaxChunks && e == cudaSuccess; c++) {
        for (int i = 0; i < 6 && e == cudaSuccess; i++) e = cudaEventCreate(&sc->cev[c][i]);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sc->chunk_done[c], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sc->copies_done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sc->ev_exchange, cudaEventDisableTiming);
    if (e != cudaSuccess) return fail(rth::cuda_fail(e, "cudaEventCreate", __FILE__, __LINE__));
    sc->n_vertices = d->n_vertices; sc->n_triangles = d->n_triangles; sc->n_meshes = d->n_meshes;
    sc->n_instances = d->n_instances; sc->n_lights = d->n_lights;
    sc->camera = d->camera; sc->env = d->env; sc->width = d->width; sc->height = d->height;
    sc->instances.assign(d->instances, d->instances + d->n_instances);
    sc->meshes.assign(d->meshes, d->meshes + d->n_meshes);

    rth::HostScene hs;
    rth::prepare_host_scene(d, hs);
    sc->any_refractive = hs.any_refractive;
    sc->any_secondary = hs.any_secondary;
    sc->cull_ok = hs.cull_ok;
    sc->n_materials = (int)hs.mats.size();
    sc->mesh_tri_count = hs.mesh_tri_count;
    sc->meshes_identity = true;
    for (const MeshRec& m : hs.meshes) if (!m.pose.identity_rot) sc->meshes_identity = false;
    int rc;
    if ((rc = upload(sc->d_verts, hs.verts, sc->stream)) || (rc = upload(sc->d_norms, hs.norms, sc->stream)) ||
        (rc = upload(sc->d_tris, hs.tris, sc->stream)) || (rc = upload(sc->d_mats, hs.mats, sc->stream)) ||
        (rc = upload(sc->d_meshes, hs.meshes, sc->stream)) || (rc = upload(sc->d_mesh_box, hs.mesh_box, sc->stream)) ||
        (rc = upload(sc->d_lights, hs.lights, sc->stream)))
        return fail(rc);
    if (sc->n_triangles > 0) {
        if (sc->d_tri_pre.reserve(sizeof(float4) * (size_t)sc->n_triangles)) return fail(RT_ERR_OUT_OF_MEMORY);
        k_tri_precompute<<<(sc->n_triangles + kThreads - 1) / kThreads, kThreads, 0, sc->stream>>>(
            sc->d_tris.as<TriRec>(), sc->n_triangles, sc->d_verts.as<float4>(), sc->n_vertices, sc->d_tri_pre.as<float4>());
        e = cudaGetLastError();
        if (e != cudaSuccess) return fail(rth::cuda_fail(e, "k_tri_precompute", __FILE__, __LINE__));
    }
    sc->textured = !hs.texc.empty();
    if (sc->textured) {
        sc->atlas_w = hs.atlas_w; sc->atlas_h = hs.atlas_h;
        if ((rc = upload(sc->d_texc, hs.texc, sc->stream)) || (rc = upload(sc->d_atlas, hs.atlas, sc->stream))) return fail(rc);
    }
    const size_t n_pinned = 64 + (size_t)C_N * RtScene::kMaxChunks;
    if (sc->d_counters.reserve(sizeof(unsigned long long) * C_N * RtScene::kMaxChunks)) return fail(RT_ERR_OUT_OF_MEMORY);
    e = cudaHostAlloc((void**)&sc->h_pinned, n_pinned * sizeof(unsigned long long), cudaHostAllocDefault);
    if (e != cudaSuccess) return fail(rth::cuda_fail(e, "cudaHostAlloc", __FILE__, __LINE__));
    std::memset(sc->h_pinned, 0, n_pinned * sizeof(unsigned long long));
    if ((rc = upload_instances(sc))) return fail(rc);
    e = cudaStreamSynchronize(sc->stream);
    if (e != cudaSuccess) return fail(rth::cuda_fail(e, "cudaStreamSynchronize", __FILE__, __LINE__));
    *out = sc;
    return RT_OK;
}

int rt_scene_destroy(RtScene* sc) {
    if (!sc) return RT_OK;
    if (sc->stream) cudaStreamSynchronize(sc->stream);
    rth::DevBuf* all[] = {&sc->d_verts, &sc->d_norms, &sc->d_tris, &sc->d_tri_pre, &sc->d_meshes, &sc->d_insts, &sc->d_mats, &sc->d_lights,
                          &sc->d_mesh_box, &sc->d_texc, &sc->d_atlas, &sc->d_push_ctr, &sc->d_sort_keys_a, &sc->d_sort_keys_b,
                          &sc->d_sort_vals_b, &sc->d_sort_hist, &sc->d_canvas, &sc->d_counters, &sc->d_hitq, &sc->d_paths, &sc->d_queue_a,
                          &sc->d_queue_b, &sc->d_ids, &sc->d_shadowq, &sc->d_occl, &sc->d_shq_count};
    for (rth::DevBuf* b : all) b->release();
    sc->bvh.release();
    if (sc->copy_stream) cudaStreamSynchronize(sc->copy_stream);
    for (int i = 0; i < 12; i++) if (sc->ev[i]) cudaEventDestroy(sc->ev[i]);
    for (int c = 0; c < RtScene::kMaxChunks; c++) {
        for (int i = 0; i < 6; i++) if (sc->cev[c][i]) cudaEventDestroy(sc->cev[c][i]);
        if (sc->chunk_done[c]) cudaEventDestroy(sc->chunk_done[c]);
    }
    for (rth::FrameGraph& g : sc->graphs) if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
    if (sc->copies_done) cudaEventDestroy(sc->copies_done);
    if (sc->ev_exchange) cudaEventDestroy(sc->ev_exchange);
    if (sc->lane_fork) cudaEventDestroy(sc->lane_fork);
    for (int c = 0; c < RtScene::kMaxChunks - 1; c++) {
        if (sc->lane_join[c]) cudaEventDestroy(sc->lane_join[c]);
        if (sc->lane_stream[c]) { cudaStreamSynchronize(sc->lane_stream[c]); cudaStreamDestroy(sc->lane_stream[c]); }
    }
    if (sc->copy_stream) cudaStreamDestroy(sc->copy_stream);
    if (sc->own_stream) cudaStreamDestroy(sc->own_stream);
    if (sc->h_pinned) cudaFreeHost(sc->h_pinned);
    if (sc->h_inst) cudaFreeHost(sc->h_inst);
    delete sc;
    return RT_OK;
}

int rt_scene_set_camera(RtScene* sc, const RtCamera* cam) {
    if (!sc || !cam) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    sc->camera = *cam;
    return RT_OK;
}
int rt_scene_set_env(RtScene* sc, const RtEnv* env) {
    if (!sc || !env) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    if (env->depth < 0 || env->depth >= RT_MAX_FRAMES) { set_error("depth must be in 0..9"); return RT_ERR_LIMIT; }
    sc->env = *env;
    return RT_OK;
}
int rt_scene_set_instances(RtScene* sc, const RtInstance* inst, int32_t n) {
    if (!sc || n < 0 || (n && !inst)) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    for (int i = 0; i < n; i++)
        if (inst[i].mesh < 0 || inst[i].mesh >= sc->n_meshes) { set_error("instance mesh index out of range"); return RT_ERR_INVALID_ARG; }
    sc->instances.assign(inst, inst + n);
    sc->n_instances = n;
    sc->inst_gen++;
    return upload_instances(sc);
}
int rt_scene_set_canvas_size(RtScene* sc, int32_t w, int32_t h) {
    if (!sc || w <= 0 || h <= 0 || w > 65535 || h > 65535) { set_error("invalid canvas size"); return RT_ERR_INVALID_ARG; }
    if ((long long)w * h > (1ll << 30)) { set_error("canvas larger than 2^30 pixels"); return RT_ERR_LIMIT; }
    sc->width = w; sc->height = h;
    for (RtScene::SbTuner& t : sc->tuners) t = RtScene::SbTuner();
    sc->camera.canvas_width = (float)w;
    sc->camera.canvas_height = (float)h;
    return RT_OK;
}

int rt_bvh_build(RtScene* sc, RtBuildStats* stats) {
    if (!sc) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
    return rth::build_bvh(sc, stats);
}
int rt_bvh_refit(RtScene* sc, RtBuildStats* stats) {
    if (!sc) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
    if (sc->pending) { if (int rc = rt_frame_finish(sc, nullptr)) return rc; }
    if (stats) { stats->n_prims = sc->bvh.n_prims; stats->n_bvh2_nodes = sc->bvh.n_bvh2_nodes; stats->n_bvh8_nodes = 0; stats->kernel_launches = 0; stats->ms_build = 0.f; }
    return rth::refit_bvh(sc, stats);
}
int rt_bvh_free(RtScene* sc) {
    if (!sc) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
    sc->bvh.release();
    return RT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------- frame driver
namespace {

template <class K>
static int persistent_grid(K kernel, int sm_count) {
    int per_sm = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    return sm_count * per_sm;
}

// Wave 0 has three forms (all give the same frame; measured in profiles/r2_summary.md):
//   split          k_primary (per-lane traversal) -> k_shadow_gen -> k_shadow -> k_resolve       RT_RENDER_SPLIT_SHADE
//   beam primary   k_tile<1> (warp-coherent closest hits, rt_beam.cuh) -> the same split shading  the default
//   tile kernels   k_tile<1> -> k_tile<2> (or the one-kernel k_tile<0>): no shadow queues         RT_RENDER_TILE_KERNEL
// Is wave 0 of this frame run by the tile kernels? Opt-in (flag, or RT_TILE_KERNEL=1 in the environment), for scenes whose
// shadow rays are occlusion queries (no refractive material), BVH traversal with the reference's leaf-box rule.
static bool tile_kernel_eligible(const RtScene* sc, const SceneView& v, int mode) {
    static const int env_on = [] { const char* e = std::getenv("RT_TILE_KERNEL"); return (e && std::atoi(e)) ? 1 : 0; }();
    return mode == TR_BVH_BOX && (env_on || sc->tile_kernel) && !sc->fused_shade && !sc->split_shade && !v.any_refractive;  // not in the tolerance mode
}
// The closest hits of the camera rays by the warp-coherent k_tile<1> in front of the split shading kernels (default).
static bool beam_primary_eligible(const RtScene* sc, int mode) {
    static const int env_off = [] { const char* e = std::getenv("RT_BEAM_PRIMARY"); return (e && !std::atoi(e)) ? 1 : 0; }();
    return (mode == TR_BVH_BOX || mode == TR_BVH_BOX_FAST) && !env_off && !sc->fused_shade && !sc->split_shade;
}

template <int MODE, bool STATS>
static int run_wavefront(RtScene* sc, const SceneView& v, const TileMap& tm, uint32_t* canvas, int* waves,
                         unsigned long long* ctr, cudaEvent_t* cev, const PeerOut* peers) {
    // ctr: this chunk's counter block; cev: its events {start, primary done, shadow rays generated, shadow rays traced,
    // wave 0 shaded, bounce waves done}. peers: k_tile stores the pixels into these peer canvases and signals them
    // (only passed when the tile kernel is eligible and the frame has no secondary rays).
    cudaStream_t st = sc->cur_stream ? sc->cur_stream : sc->stream;  // a lane's stream (render_enqueue) or the scene's
    const size_t q_off = (size_t)sc->cur_off;  // this lane's first slot of the hit / shadow queues (0 without lanes)
    RT_CUDA(rth::record(sc, cev[0], st));
    // tolerance mode: only the default pipeline (k_tile<1>, k_shadow_beam / k_shadow, k_bounce_waves) has a fast variant; the
    // other branches run their exact kernels (BM), which is always an acceptable answer
    constexpr int BM = MODE == TR_BVH_BOX_FAST ? TR_BVH_BOX : MODE;
    long long n_pix = (long long)tm.n_work() * 32;
    *waves = 1;
    sc->split_shade_used = false;
    sc->tile_kernel_used = false;
    int path_stride = (int)(sizeof(PathHead) + sizeof(Frame) * (size_t)(v.depth + 1));
    const bool secondary = v.any_secondary && v.depth > 0;
    const bool loop = v.any_refractive != 0;  // shadow rays walk through refractive occluders
    unsigned long long n_hits = 0;
    size_t hit_cap = (size_t)std::max<long long>(n_pix, 1);
    // k_tile starts the paths of reflective hits itself, so their stacks are allocated for every pixel of the chunk (only
    // the slots of real paths are ever touched); beyond 32 GB the split pipeline, which can size them after k_primary, runs
    bool use_tile = tile_kernel_eligible(sc, v, MODE) && !(secondary && hit_cap * (size_t)path_stride > ((size_t)32 << 30));
    if (peers && !(use_tile && !secondary)) { set_error("internal: peer stores need the tile kernel"); return RT_ERR_INVALID_ARG; }
    if (!use_tile || secondary) {
        if (sc->d_hitq.reserve(sizeof(HitRec) * (q_off + (size_t)std::max<long long>(n_pix, 1)))) return RT_ERR_OUT_OF_MEMORY;
    }
    sc->wave0_form = use_tile ? 1 : 0;
    if (!use_tile && beam_primary_eligible(sc, MODE)) {
        PeerOut po0;
        std::memset(&po0, 0, sizeof po0);
        constexpr bool FAST = MODE == TR_BVH_BOX_FAST;
        static const int g0 = persistent_grid(k_tile<STATS, false, false, 1, FAST>, sc->sm_count);
        k_tile<STATS, false, false, 1, FAST><<<g0, kThreads, 0, st>>>(v, tm, ctr, canvas, po0, nullptr, 0, nullptr, sc->d_hitq.as<HitRec>() + q_off);
        sc->kernel_launches++;
        sc->wave0_form = 2;
        RT_CUDA(rth::record(sc, cev[1], st));
    } else if (!use_tile) {
        int g0 = persistent_grid(k_primary<BM, STATS, false>, sc->sm_count);
        static const int aligned_hitq = [] { const char* e = std::getenv("RT_PACKED_HITQ"); return (e && std::atoi(e)) ? 0 : 1; }();
        k_primary<BM, STATS, false><<<g0, kThreads, 0, st>>>(v, tm, ctr, sc->d_hitq.as<HitRec>() + q_off, canvas, nullptr, nullptr, nullptr, aligned_hitq);
        sc->kernel_launches++;
        RT_CUDA(rth::record(sc, cev[1], st));
    }
    if (secondary) {
        // Path stacks for every pixel when that is affordable (no host round trip after k_primary);
        // otherwise they are allocated for the hits actually found (one 8-byte read-back).
        if (!use_tile && hit_cap * (size_t)path_stride > ((size_t)4 << 30)) {
            RT_CUDA(cudaMemcpyAsync(&n_hits, ctr + C_HITQ, 8, cudaMemcpyDeviceToHost, st));
            RT_CUDA(cudaStreamSynchronize(st));
            hit_cap = (size_t)std::max<unsigned long long>(n_hits & 0xffffffffull, 1);  // queue slots (>= hits)
        }
        if (sc->d_paths.reserve(hit_cap * path_stride) || sc->d_queue_a.reserve(4 * hit_cap) || sc->d_queue_b.reserve(4 * hit_cap))
            return RT_ERR_OUT_OF_MEMORY;
    }
    unsigned char* occl = nullptr;
    float4* att4 = nullptr;
    ShadowRec* shq = nullptr;
    if (!sc->fused_shade && (!use_tile || secondary)) {
        const size_t nl = (size_t)std::max(v.n_lights, 1);
        size_t cap = (q_off + hit_cap) * nl;
        if (sc->d_shadowq.reserve(sizeof(ShadowRec) * cap) || sc->d_occl.reserve((loop ? sizeof(float4) : 1) * cap) ||
            sc->d_shq_count.reserve(sizeof(unsigned) * nl * (size_t)(sc->cur_lane + 1)))
            return RT_ERR_OUT_OF_MEMORY;
        shq = sc->d_shadowq.as<ShadowRec>() + q_off * nl;
        if (loop) att4 = sc->d_occl.as<float4>() + q_off * nl; else occl = sc->d_occl.as<unsigned char>() + q_off * nl;
    }
    unsigned* shq_count = sc->d_shq_count.as<unsigned>() + (size_t)std::max(v.n_lights, 1) * (size_t)sc->cur_lane;
    // one any-hit / closest-loop traversal of the shadow rays queued in shq
    auto trace_shadows = [&](int ctr_n) {
        static const int static_split = [] { const char* e = std::getenv("RT_SHADOW_STATIC"); return (e && std::atoi(e)) ? 1 : 0; }();
        unsigned long long* work = static_split ? nullptr : ctr + C_WORK1;  // zeroed with the frame's counters
        // Wave 0 of an opaque scene behind k_tile<1>: the queues are tile-aligned, so one warp-coherent occlusion walk per
        // tile and light is possible (k_shadow_beam). Which kernel wins depends on the scene (world16 at 4K: 0.61 ms against
        // 0.70 per lane; world8_stress: 0.40 against 0.31, where most shadow rays are blocked at once and a lane-wise walk
        // stops early), and so does the order the beam kernel takes the queue in: front to back wins on the whole 4K frame
        // (0.606 ms against 0.658), back to front (the image's lower rows, its longest units, first) on an eighth of it
        // (0.133 against 0.162). A scene therefore times the three on its first six frames of a tile shape (same flags
        // each time) and keeps the fastest; RT_BEAM_SHADOW=0|1|2 forces lane-wise / beam / beam from the back.
        static const int beam_env = [] { const char* e = std::getenv("RT_BEAM_SHADOW"); return e ? std::min(std::max(std::atoi(e), 0), 2) : -1; }();
        const bool beam_possible = !loop && sc->wave0_form == 2 && ctr_n == C_HITQ && v.n_lights > 0;
        const RtScene::SbTuner& tuner = sc->tuners[sc->sb_cur];
        const int beam_pick = beam_env >= 0 ? beam_env : tuner.choice >= 0 ? tuner.choice : (tuner.frames + 1) % 3;
        sc->sb_used = beam_possible ? beam_pick : -1;
        if (beam_possible && beam_pick) {
            constexpr bool FAST = MODE == TR_BVH_BOX_FAST;
            static const int g = persistent_grid(k_shadow_beam<STATS, FAST>, sc->sm_count);
            k_shadow_beam<STATS, FAST><<<g, kThreads, 0, st>>>(v, ctr, ctr_n, shq, occl, ctr + C_WORK1, beam_pick == 2);
            sc->kernel_launches++;
            return;
        }
        static const int lane_back_env = [] { const char* e = std::getenv("RT_SHADOW_LANE_ORDER"); return (e && std::atoi(e)) ? 1 : 0; }();
        const int lane_back = (beam_possible && lane_back_env && !static_split) ? 1 : 0;
        if (loop) k_shadow<MODE, STATS, true><<<persistent_grid(k_shadow<MODE, STATS, true>, sc->sm_count), kThreads, 0, st>>>(
            v, ctr, ctr_n, shq, shq_count, occl, att4, work, 0);
        else k_shadow<MODE, STATS, false><<<persistent_grid(k_shadow<MODE, STATS, false>, sc->sm_count), kThreads, 0, st>>>(
            v, ctr, ctr_n, shq, shq_count, occl, att4, work, lane_back);
        sc->kernel_launches++;
    };
    HitRec* hitq = sc->d_hitq.as<HitRec>() + q_off;
    unsigned char* paths = sc->d_paths.as<unsigned char>();
    if (use_tile) {
        PeerOut po;
        std::memset(&po, 0, sizeof po);
        if (peers) po = *peers;
        int* qa = secondary ? sc->d_queue_a.as<int>() : nullptr;
        // RT_TILE_PHASES=1: the whole tile in one kernel; 2 (default): closest hits, then shading, through the hit queue
        static const int phases = [] { const char* e = std::getenv("RT_TILE_PHASES"); int v = e ? std::atoi(e) : 2; return v == 1 ? 1 : 2; }();
        HitRec* hq = nullptr;
        if (phases == 2) {
            if (sc->d_hitq.reserve(sizeof(HitRec) * (size_t)std::max<long long>(n_pix, 1))) return RT_ERR_OUT_OF_MEMORY;
            hq = sc->d_hitq.as<HitRec>();
        }
#define RT_LAUNCH_TILE(SEC, PEER, PH)                                                                              \
        {                                                                                                          \
            static const int g = persistent_grid(k_tile<STATS, SEC, PEER, PH>, sc->sm_count);                      \
            k_tile<STATS, SEC, PEER, PH><<<g, kThreads, 0, st>>>(v, tm, ctr, canvas, po, paths, path_stride, qa, hq); \
            sc->kernel_launches++;                                                                                 \
        }
        if (phases == 2) {
            if (peers) { RT_LAUNCH_TILE(false, true, 1) RT_LAUNCH_TILE(false, true, 2) }
            else if (secondary) { RT_LAUNCH_TILE(true, false, 1) RT_LAUNCH_TILE(true, false, 2) }
            else { RT_LAUNCH_TILE(false, false, 1) RT_LAUNCH_TILE(false, false, 2) }
        } else {
            if (peers) RT_LAUNCH_TILE(false, true, 0)
            else if (secondary) RT_LAUNCH_TILE(true, false, 0)
            else RT_LAUNCH_TILE(false, false, 0)
        }
#undef RT_LAUNCH_TILE
        sc->kernel_launches--;  // the common tail below counts wave 0's last kernel
        sc->tile_kernel_used = true;
        RT_CUDA(rth::record(sc, cev[1], st));
        RT_CUDA(rth::record(sc, cev[2], st));
        RT_CUDA(rth::record(sc, cev[3], st));
    } else if (!sc->fused_shade) {
        k_shadow_gen<false><<<persistent_grid(k_shadow_gen<false>, sc->sm_count), kThreads, 0, st>>>(v, ctr, C_HITQ, hitq, shq, shq_count, occl, att4, nullptr, 0);
        sc->kernel_launches++;
        RT_CUDA(rth::record(sc, cev[2], st));
        trace_shadows(C_HITQ);
        RT_CUDA(rth::record(sc, cev[3], st));
        if (secondary)
            k_resolve<true><<<persistent_grid(k_resolve<true>, sc->sm_count), kThreads, 0, st>>>(
                v, ctr, hitq, occl, att4, canvas, paths, path_stride, sc->d_queue_a.as<int>());
        else
            k_resolve<false><<<persistent_grid(k_resolve<false>, sc->sm_count), kThreads, 0, st>>>(
                v, ctr, hitq, occl, att4, canvas, nullptr, 0, nullptr);
        sc->split_shade_used = true;
    } else if (!v.any_secondary) {
        int g1 = persistent_grid(k_shade<BM, STATS, true>, sc->sm_count);
        k_shade<BM, STATS, true><<<g1, kThreads, 0, st>>>(v, ctr, hitq, canvas, nullptr, 0, nullptr);
    } else {
        int g1 = persistent_grid(k_shade<BM, STATS, false>, sc->sm_count);
        k_shade<BM, STATS, false><<<g1, kThreads, 0, st>>>(v, ctr, hitq, canvas, paths, path_stride, sc->d_queue_a.as<int>());
    }
    sc->kernel_launches++;
    RT_CUDA(rth::record(sc, cev[4], st));
    if (secondary) {
        int in = C_QUEUE_A, outc = C_QUEUE_B;
        int* qin = sc->d_queue_a.as<int>();
        int* qout = sc->d_queue_b.as<int>();
        if (sc->fused_shade) {
            for (;;) {
                unsigned long long live = 0;
                RT_CUDA(cudaMemcpyAsync(&live, ctr + in, 8, cudaMemcpyDeviceToHost, st));
                RT_CUDA(cudaStreamSynchronize(st));
                if (live == 0) break;
                RT_CUDA(cudaMemsetAsync(ctr + outc, 0, 8, st));
                k_bounce<BM, STATS><<<persistent_grid(k_bounce<BM, STATS>, sc->sm_count), kThreads, 0, st>>>(
                    v, ctr, in, outc, qin, qout, canvas, paths, path_stride);
                sc->kernel_launches++;
                (*waves)++;
                std::swap(in, outc);
                std::swap(qin, qout);
                if (*waves > 4096) { set_error("bounce waves did not terminate"); return RT_ERR_LIMIT; }
            }
        } else if (sc->sort_rays) {
            // RT_RENDER_SORT_RAYS: one launch per phase and wave, the live queue sorted by origin cell + direction octant
            // before each trace phase (one 8-byte read-back per wave tells the host how many paths are alive)
            if (!sc->bvh.scene_bounds.p) { set_error("no scene bounds to sort rays by"); return RT_ERR_NO_BVH; }
            const float* bounds6 = (const float*)((const char*)sc->bvh.scene_bounds.p + 32);
            for (;;) {
                unsigned long long live = 0;
                RT_CUDA(cudaMemcpyAsync(&live, ctr + in, 8, cudaMemcpyDeviceToHost, st));
                RT_CUDA(cudaStreamSynchronize(st));
                if (live == 0) break;
                const int n = (int)live;
                if (sc->d_sort_keys_a.reserve(8 * (size_t)n) || sc->d_sort_keys_b.reserve(8 * (size_t)n) ||
                    sc->d_sort_vals_b.reserve(4 * (size_t)n) || sc->d_sort_hist.reserve(rth::radix_hist_bytes(n)))
                    return RT_ERR_OUT_OF_MEMORY;
                unsigned long long* ka = sc->d_sort_keys_a.as<unsigned long long>();
                unsigned long long* kb = sc->d_sort_keys_b.as<unsigned long long>();
                int* va = qout;  // the other queue is free until this wave's advance phase: sort into it
                int* vb = sc->d_sort_vals_b.as<int>();
                k_path_keys<<<(n + kThreads - 1) / kThreads, kThreads, 0, st>>>(qin, n, paths, path_stride, bounds6, ka, va);
                sc->kernel_launches += 1 + rth::radix_sort_pairs(st, ka, kb, va, vb, sc->d_sort_hist.as<unsigned>(), n, 3);
                // three passes: the sorted slots are in what `vb` pointed at first, i.e. d_sort_vals_b; copy them into the queue
                RT_CUDA(cudaMemcpyAsync(qin, va, 4 * (size_t)n, cudaMemcpyDeviceToDevice, st));
                RT_CUDA(cudaMemsetAsync(ctr + outc, 0, 8, st));
                RT_CUDA(cudaMemsetAsync(shq_count, 0, sizeof(unsigned) * (size_t)std::max(v.n_lights, 1), st));
                k_wave_trace<BM, STATS><<<persistent_grid(k_wave_trace<BM, STATS>, sc->sm_count), kThreads, 0, st>>>(
                    v, ctr, in, qin, paths, path_stride, hitq);
                k_shadow_gen<true><<<persistent_grid(k_shadow_gen<true>, sc->sm_count), kThreads, 0, st>>>(
                    v, ctr, in, hitq, shq, shq_count, occl, att4, paths, path_stride);
                if (loop) k_shadow<MODE, STATS, true><<<persistent_grid(k_shadow<MODE, STATS, true>, sc->sm_count), kThreads, 0, st>>>(
                    v, ctr, in, shq, shq_count, occl, att4, nullptr, 0);
                else k_shadow<MODE, STATS, false><<<persistent_grid(k_shadow<MODE, STATS, false>, sc->sm_count), kThreads, 0, st>>>(
                    v, ctr, in, shq, shq_count, occl, att4, nullptr, 0);
                k_wave_advance<<<persistent_grid(k_wave_advance, sc->sm_count), kThreads, 0, st>>>(
                    v, ctr, in, outc, hitq, occl, att4, qout, canvas, paths, path_stride);
                sc->kernel_launches += 4;
                (*waves)++;
                std::swap(in, outc);
                std::swap(qin, qout);
                if (*waves > 4096) { set_error("bounce waves did not terminate"); return RT_ERR_LIMIT; }
            }
        } else {
            // every wave inside one cooperative launch; the wave count comes back with the frame counters
            WaveArgs wa;
            wa.ctr = ctr; wa.queue_a = qin; wa.queue_b = qout; wa.whit = hitq; wa.shq = shq; wa.shq_count = shq_count;
            wa.occl = occl; wa.att4 = att4; wa.canvas = canvas; wa.paths = paths; wa.path_stride = path_stride;
            wa.max_waves = 4096;
            wa.barrier = (unsigned*)(ctr + C_BARRIER);  // zeroed with the counters at the start of the frame
            SceneView vv = v;
            void* args[] = {&vv, &wa};
            if (loop) {
                static const int g = persistent_grid(k_bounce_waves<MODE, STATS, true>, sc->sm_count);
                RT_CUDA(cudaLaunchCooperativeKernel((void*)k_bounce_waves<MODE, STATS, true>, dim3(g), dim3(kThreads), args, 0, st));
            } else {
                static const int g = persistent_grid(k_bounce_waves<MODE, STATS, false>, sc->sm_count);
                RT_CUDA(cudaLaunchCooperativeKernel((void*)k_bounce_waves<MODE, STATS, false>, dim3(g), dim3(kThreads), args, 0, st));
            }
            sc->kernel_launches++;
            *waves = -1;  // 1 + ctr[C_WAVES], filled in by the caller after the counter read-back
        }
    }
    RT_CUDA(rth::record(sc, cev[5], st));
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

template <int MODE>
static int run_mega(RtScene* sc, const SceneView& v, const TileMap& tm, uint32_t* canvas, int semantics) {
    unsigned long long* ctr = sc->d_counters.as<unsigned long long>();
    if (semantics == RT_SEMANTICS_CPU) {
        // propagate_helper recursion: RT_CPU_SEM_MAX_LEVEL frames of < 1 KB each
        RT_CUDA(cudaDeviceSetLimit(cudaLimitStackSize, 32 * 1024));
        int g = persistent_grid(k_mega<MODE, RT_SEMANTICS_CPU>, sc->sm_count);
        k_mega<MODE, RT_SEMANTICS_CPU><<<g, kThreads, 0, sc->stream>>>(v, tm, ctr, canvas);
        sc->kernel_launches++;
        RT_CUDA(cudaGetLastError());
        RT_CUDA(cudaStreamSynchronize(sc->stream));
        RT_CUDA(cudaDeviceSetLimit(cudaLimitStackSize, 2048));
        return RT_OK;
    } else {
        int g = persistent_grid(k_mega<MODE, RT_SEMANTICS_GPU>, sc->sm_count);
        k_mega<MODE, RT_SEMANTICS_GPU><<<g, kThreads, 0, sc->stream>>>(v, tm, ctr, canvas);
    }
    sc->kernel_launches++;
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

// Device canvas -> host canvas for the owned tiles [k0, k0 + kn), whose pixels are whole rows of the host canvas.
// bands == false: one rank, row-major device canvas; the tiles are tile rows [k0 / tiles_x, ...), i.e. one contiguous
// block of rows. bands == true: `dcanvas` is the rank's slab of full-width bands (pitch tile_w) and `host` the row-major
// W x H canvas every rank scatters its bands into: band k sits at rows (k world + rank) tile_h ..., one strided 3-D copy.
static int copy_rows_to_host(RtScene* sc, const TileMap& tm, int k0, int kn, const uint32_t* dcanvas, uint32_t* host, bool bands,
                             cudaStream_t st) {
    (void)sc;
    if (kn <= 0) return RT_OK;
    if (!bands) {
        const long long y0 = (long long)(k0 / tm.tiles_x) * tm.tile_h;
        const long long y1 = std::min<long long>((long long)((k0 + kn) / tm.tiles_x) * tm.tile_h, tm.H);
        if (y1 > y0) RT_CUDA(cudaMemcpyAsync(host + y0 * tm.W, dcanvas + y0 * tm.W, 4 * (size_t)(y1 - y0) * tm.W, cudaMemcpyDeviceToHost, st));
        return RT_OK;
    }
    // the frame's last band may be cut by the canvas edge: copy it on its own
    int full = kn;
    const long long last_T = (long long)(k0 + kn - 1) * tm.world + tm.rank;
    const long long last_rows = std::min<long long>(tm.tile_h, tm.H - last_T * tm.tile_h);
    if (last_rows < tm.tile_h) full--;
    if (full > 0) {
        cudaMemcpy3DParms c;
        std::memset(&c, 0, sizeof c);
        c.srcPtr = make_cudaPitchedPtr((void*)(dcanvas + (size_t)k0 * tm.tile_h * tm.tile_w), 4 * (size_t)tm.tile_w, 4 * (size_t)tm.tile_w, tm.tile_h);
        c.dstPtr = make_cudaPitchedPtr((void*)(host + ((size_t)k0 * tm.world + tm.rank) * tm.tile_h * tm.W), 4 * (size_t)tm.W, 4 * (size_t)tm.W,
                                       (size_t)tm.world * tm.tile_h);
        c.extent = make_cudaExtent(4 * (size_t)tm.W, tm.tile_h, full);
        c.kind = cudaMemcpyDeviceToHost;
        RT_CUDA(cudaMemcpy3DAsync(&c, st));
    }
    if (full < kn && last_rows > 0)
        RT_CUDA(cudaMemcpy2DAsync(host + (size_t)last_T * tm.tile_h * tm.W, 4 * (size_t)tm.W,
                                  dcanvas + (size_t)(k0 + kn - 1) * tm.tile_h * tm.tile_w, 4 * (size_t)tm.tile_w, 4 * (size_t)tm.W,
                                  (size_t)last_rows, cudaMemcpyDeviceToHost, st));
    return RT_OK;
}

// built == nullptr: blocking build. Otherwise the build is only enqueued and *built tells the
// caller to run rth::finish_build() after its own stream synchronisation.
static int ensure_bvh(RtScene* sc, bool rebuild, bool* built, bool refit = false) {
    if (built) *built = false;
    if (sc->bvh.valid && !rebuild && !refit) return RT_OK;
    if (!rebuild && rth::refit_possible(sc)) {  // instances moved (or RT_RENDER_REFIT_BVH): keep the tree
        if (int rc = rth::refit_bvh(sc, nullptr, built == nullptr)) return rc;
        if (built) *built = true;
        return RT_OK;
    }
    if (int rc = rth::build_bvh(sc, nullptr, built == nullptr)) return rc;
    if (built) *built = true;
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_tiled_slab_pixels(RtScene* sc, const RtRenderParams* p, int64_t* out) {
    if (!sc || !out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    TileMap tm;
    if (int rc = make_tilemap(sc, p, tm)) return rc;
    *out = slab_pixels(tm);
    return RT_OK;
}

int rt_tile_map(int32_t width, int32_t height, int32_t tile_w, int32_t tile_h, int32_t rank, int32_t world,
                int64_t* slab_index, int64_t* slab_pix) {
    if (width <= 0 || height <= 0 || !slab_index) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    RtScene dims;  // only width/height are read
    dims.width = width; dims.height = height;
    RtRenderParams p;
    std::memset(&p, 0, sizeof p);
    p.flags = RT_RENDER_TILED_OUTPUT; p.rank = rank; p.world = world; p.tile_w = tile_w; p.tile_h = tile_h;
    TileMap tm;
    if (int rc = make_tilemap(&dims, &p, tm)) return rc;
    for (long long i = 0; i < (long long)width * height; i++) slab_index[i] = -1;
    for (int q = 0; q < tm.n_work(); q++)
        for (int lane = 0; lane < 32; lane++) {
            int x, y, out;
            if (tm.locate(q, lane, x, y, out)) slab_index[(long long)y * width + x] = out;
        }
    if (slab_pix) *slab_pix = slab_pixels(tm);
    return RT_OK;
}

int rt_untile(RtScene* sc, const RtRenderParams* p, const uint32_t* gathered, uint32_t* canvas) {
    if (!sc || !gathered || !canvas) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    TileMap tm;
    if (int rc = make_tilemap(sc, p, tm)) return rc;
    long long n = (long long)tm.W * tm.H;
    k_untile<<<(unsigned)((n + kThreads - 1) / kThreads), kThreads, 0, sc->stream>>>(tm, slab_pixels(tm), gathered, canvas);
    sc->kernel_launches++;
    RT_CUDA(cudaGetLastError());
    if (!(p && (p->flags & RT_RENDER_ASYNC))) RT_CUDA(cudaStreamSynchronize(sc->stream));
    return RT_OK;
}

// Time-out of the mailbox wait: 2 s by default (RT_MAILBOX_TIMEOUT_MS overrides it), in SM cycles at ~2 GHz.
static long long mailbox_timeout_cycles() {
    static const long long cyc = [] {
        const char* e = std::getenv("RT_MAILBOX_TIMEOUT_MS");
        double ms = e ? std::atof(e) : 2000.0;
        if (!(ms > 0.0)) ms = 2000.0;
        return (long long)(ms * 2.0e6);
    }();
    return cyc;
}
// The word k_wait_mailbox / the fused exchange report a time-out in: h_pinned[48] (page-locked, device-addressable under UVA).
static unsigned* exchange_status(RtScene* sc) { return (unsigned*)(sc->h_pinned + 48); }
// After a stream synchronisation: a mailbox wait of this scene that gave up makes the frame an error, not a silent hole.
static int check_exchange(RtScene* sc) {
    unsigned* st = exchange_status(sc);
    const unsigned missing = *(volatile unsigned*)st;
    if (!missing) return RT_OK;
    *st = 0u;
    char buf[160];
    std::snprintf(buf, sizeof buf, "peer exchange timed out: ranks with mask 0x%x never signalled (or, gathering rank, never acknowledged) this frame", missing);
    set_error(buf);
    return RT_ERR_CUDA;
}

static int push_tiles_impl(RtScene* sc, const RtRenderParams* p, const uint32_t* slab, uint32_t* const* peer_canvases, int32_t n_peers,
                           int64_t canvas_pixels, uint32_t epoch, bool signal) {
    if (!sc || !slab || !peer_canvases || n_peers < 1 || n_peers > RT_MAX_PEERS) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    TileMap tm;
    if (int rc = make_tilemap(sc, p, tm)) return rc;
    if (tm.tile_w % 4) { set_error("tile_w must be a multiple of 4"); return RT_ERR_INVALID_ARG; }
    if (signal && (n_peers != tm.world || canvas_pixels < (int64_t)tm.W * tm.H)) {
        set_error("signalled push needs one canvas per rank and the canvas size the mailbox sits behind");
        return RT_ERR_INVALID_ARG;
    }
    PeerCanvases pc;
    pc.n = n_peers;
    for (int r = 0; r < n_peers; r++) {
        if (!peer_canvases[r]) { set_error("null peer canvas"); return RT_ERR_INVALID_ARG; }
        pc.p[r] = peer_canvases[r];
    }
    PushSignal sg;
    sg.mailbox_offset = canvas_pixels; sg.epoch = epoch; sg.done_blocks = nullptr; sg.epoch_ptr = nullptr;
    long long n_quads = (long long)tm.n_owned * tm.tile_h * (tm.tile_w / 4);
    if (signal) {
        if (!sc->d_push_ctr.p) {
            if (sc->d_push_ctr.reserve(16)) return RT_ERR_OUT_OF_MEMORY;
            RT_CUDA(cudaMemsetAsync(sc->d_push_ctr.p, 0, 16, sc->stream));
        }
        sg.done_blocks = sc->d_push_ctr.as<unsigned>();
        // a rank that owns no tile still has to signal: one block, no quads
        unsigned grid = (unsigned)std::max<long long>((n_quads + kThreads - 1) / kThreads, 1);
        k_push_tiles<true><<<grid, kThreads, 0, sc->stream>>>(tm, slab, pc, sg);
        PeerCanvases no_ack;
        std::memset(&no_ack, 0, sizeof no_ack);
        k_wait_mailbox<<<1, 32, 0, sc->stream>>>(peer_canvases[tm.rank], canvas_pixels, n_peers, epoch, mailbox_timeout_cycles(),
                                                 exchange_status(sc), nullptr, no_ack, 0);
        RT_CUDA(cudaEventRecord(sc->ev_exchange, sc->stream));
        sc->exchange_pending = true;
        sc->kernel_launches += 2;
    } else if (n_quads > 0) {
        k_push_tiles<false><<<(unsigned)((n_quads + kThreads - 1) / kThreads), kThreads, 0, sc->stream>>>(tm, slab, pc, sg);
        sc->kernel_launches++;
    }
    RT_CUDA(cudaGetLastError());
    if (!(p && (p->flags & RT_RENDER_ASYNC))) {
        RT_CUDA(cudaStreamSynchronize(sc->stream));
        sc->exchange_pending = false;
        return check_exchange(sc);
    }
    return RT_OK;
}

int rt_push_tiles(RtScene* sc, const RtRenderParams* p, const uint32_t* slab, uint32_t* const* peer_canvases, int32_t n_peers) {
    return push_tiles_impl(sc, p, slab, peer_canvases, n_peers, 0, 0, false);
}

int rt_push_tiles_signal(RtScene* sc, const RtRenderParams* p, const uint32_t* slab, uint32_t* const* peer_canvases, int32_t n_peers,
                         int64_t canvas_pixels, uint32_t epoch) {
    return push_tiles_impl(sc, p, slab, peer_canvases, n_peers, canvas_pixels, epoch, true);
}

// Canvas memory other processes can map: plain cudaMalloc (torch's caching allocator hands out
// sub-allocations, which CUDA IPC cannot export), exported / imported as a 64-byte CUDA IPC handle.
int rt_peer_canvas_alloc(int64_t pixels, uint32_t** dev_ptr, unsigned char handle[64]) {
    if (pixels <= 0 || !dev_ptr || !handle) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    void* p = nullptr;
    const size_t bytes = 4 * ((size_t)pixels + RT_MAX_PEERS);  // + the mailbox of rt_push_tiles_signal
    RT_CUDA(cudaMalloc(&p, bytes));
    RT_CUDA(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return rth::cuda_fail(e, "cudaIpcGetMemHandle", __FILE__, __LINE__); }
    std::memcpy(handle, &h, 64);
    *dev_ptr = (uint32_t*)p;
    return RT_OK;
}
int rt_peer_canvas_free(uint32_t* dev_ptr) {
    if (dev_ptr) RT_CUDA(cudaFree(dev_ptr));
    return RT_OK;
}
int rt_peer_canvas_open(const unsigned char handle[64], uint32_t** dev_ptr) {
    if (!handle || !dev_ptr) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    void* p = nullptr;
    RT_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *dev_ptr = (uint32_t*)p;
    return RT_OK;
}
int rt_peer_canvas_close(uint32_t* dev_ptr) {
    if (dev_ptr) RT_CUDA(cudaIpcCloseMemHandle(dev_ptr));
    return RT_OK;
}
// One host canvas for all ranks of a box (RT_RENDER_HOST_BANDS): POSIX shared memory, page-locked in every process that
// maps it, so each GPU copies its bands into it over its own PCIe link. shm_name == NULL: private pinned memory.
int rt_host_canvas_open(const char* shm_name, int64_t pixels, int32_t create, uint32_t** host_ptr) {
    if (pixels <= 0 || !host_ptr) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    *host_ptr = nullptr;
    const size_t bytes = 4 * (size_t)pixels;
    if (!shm_name) {
        void* p = nullptr;
        RT_CUDA(cudaHostAlloc(&p, bytes, cudaHostAllocPortable));
        std::memset(p, 0, bytes);
        *host_ptr = (uint32_t*)p;
        return RT_OK;
    }
    int fd = shm_open(shm_name, create ? (O_CREAT | O_RDWR) : O_RDWR, 0600);
    if (fd < 0) { set_error(std::string("shm_open failed for ") + shm_name); return RT_ERR_INVALID_ARG; }
    if (create && ftruncate(fd, (off_t)bytes) != 0) { close(fd); set_error("ftruncate failed"); return RT_ERR_OUT_OF_MEMORY; }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) { set_error("mmap of the shared canvas failed"); return RT_ERR_OUT_OF_MEMORY; }
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) { munmap(p, bytes); return rth::cuda_fail(e, "cudaHostRegister", __FILE__, __LINE__); }
    *host_ptr = (uint32_t*)p;
    return RT_OK;
}
int rt_host_canvas_close(uint32_t* host_ptr, int64_t pixels, const char* shm_name, int32_t unlink_name) {
    if (!host_ptr) return RT_OK;
    if (!shm_name) { RT_CUDA(cudaFreeHost(host_ptr)); return RT_OK; }
    cudaHostUnregister(host_ptr);
    munmap(host_ptr, 4 * (size_t)pixels);
    if (unlink_name) shm_unlink(shm_name);
    return RT_OK;
}
int rt_copy_to_host(RtScene* sc, const uint32_t* dev_ptr, uint32_t* host_ptr, int64_t pixels) {
    if (!sc || !dev_ptr || !host_ptr || pixels < 0) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    RT_CUDA(cudaMemcpyAsync(host_ptr, dev_ptr, 4 * (size_t)pixels, cudaMemcpyDeviceToHost, sc->stream));
    RT_CUDA(cudaStreamSynchronize(sc->stream));
    return RT_OK;
}

int rt_scene_set_peers(RtScene* sc, uint32_t* const* canvases, int32_t n_ranks, int64_t canvas_pixels, int32_t gather_root) {
    if (!sc) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    if (sc->pending) { if (int rc = rt_frame_finish(sc, nullptr)) return rc; }
    if (n_ranks == 0) { sc->peer_canv.clear(); sc->peer_n = 0; return RT_OK; }
    if (!canvases || n_ranks < 1 || n_ranks > RT_MAX_PEERS || canvas_pixels < (int64_t)sc->width * sc->height || gather_root >= n_ranks) {
        set_error("rt_scene_set_peers: 1..16 ranks, 2 * n_ranks canvases of at least width * height pixels, gather_root < n_ranks");
        return RT_ERR_INVALID_ARG;
    }
    for (int i = 0; i < 2 * n_ranks; i++) if (!canvases[i]) { set_error("null peer canvas"); return RT_ERR_INVALID_ARG; }
    sc->peer_canv.assign(canvases, canvases + 2 * n_ranks);
    sc->peer_n = n_ranks;
    sc->peer_pixels = canvas_pixels;
    sc->peer_root = gather_root < 0 ? -1 : gather_root;
    return RT_OK;
}
int rt_peer_frame_canvas(RtScene* sc, int32_t rank, uint32_t** canvas, uint32_t* epoch) {
    if (!sc || !canvas || sc->peer_n < 1 || rank < 0 || rank >= sc->peer_n) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    *canvas = sc->peer_canv[(size_t)(sc->peer_epoch & 1u) * sc->peer_n + rank];
    if (epoch) *epoch = sc->peer_epoch;
    return RT_OK;
}

int rt_scene_set_stream(RtScene* sc, void* stream, int32_t external) {
    if (!sc) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    if (sc->pending) {
        if (int rc = rt_frame_finish(sc, nullptr)) return rc;
    }
    RT_CUDA(cudaStreamSynchronize(sc->stream));
    sc->stream = external ? (cudaStream_t)stream : sc->own_stream;  // external handle 0 = the legacy default stream
    return RT_OK;
}

// RT_FRAME_LANES=n (1..8, default 1): how many concurrent lanes a device-resident frame without secondary rays is cut into.
static int frame_lanes_env() {
    static const int n = [] { const char* e = std::getenv("RT_FRAME_LANES"); int v = e ? std::atoi(e) : 1; return std::min(std::max(v, 1), (int)RtScene::kMaxChunks); }();
    return n;
}
static int ensure_lane_streams(RtScene* sc, int n_lanes) {
    if (!sc->lane_fork) RT_CUDA(cudaEventCreateWithFlags(&sc->lane_fork, cudaEventDisableTiming));
    for (int c = 1; c < n_lanes; c++) {
        if (!sc->lane_stream[c - 1]) RT_CUDA(cudaStreamCreateWithFlags(&sc->lane_stream[c - 1], cudaStreamNonBlocking));
        if (!sc->lane_join[c - 1]) RT_CUDA(cudaEventCreateWithFlags(&sc->lane_join[c - 1], cudaEventDisableTiming));
    }
    return RT_OK;
}

int rt_scene_set_lanes(RtScene* sc, int32_t lanes) {
    if (!sc || lanes < 0 || lanes > RtScene::kMaxChunks) { set_error("lanes must be in 0..8"); return RT_ERR_INVALID_ARG; }
    sc->lanes_req = lanes;
    return RT_OK;
}

// Enqueues one frame on the scene's stream (plain launches, or into a stream capture when sc->capturing) and leaves it
// pending; rt_render decides between this, a capture and a replay, and finishes the frame.
static int render_enqueue(RtScene* sc, const RtRenderParams* params, uint32_t* canvas) {
    RtRenderParams p;
    std::memset(&p, 0, sizeof p);
    if (params) p = *params;
    const bool peer_store = (p.flags & RT_RENDER_PEER_STORE) != 0;
    if (!canvas && !peer_store) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    if (p.semantics != RT_SEMANTICS_GPU && p.semantics != RT_SEMANTICS_CPU) { set_error("unknown semantics"); return RT_ERR_INVALID_ARG; }
    if (sc->pending) {  // an asynchronous frame was never finished: its counters are about to be reused
        if (int rc = rt_frame_finish(sc, nullptr)) return rc;
    }
    TileMap tm;
    if (int rc = make_tilemap(sc, &p, tm)) return rc;
    cudaStream_t st = sc->stream;
    sc->cur_stream = nullptr; sc->cur_off = 0; sc->cur_lane = 0;
    int launches0 = sc->kernel_launches;
    const bool no_bvh = (p.flags & RT_RENDER_NO_BVH) != 0;
    const bool want_stats = (p.flags & RT_RENDER_STATS) != 0;
    sc->fused_shade = (p.flags & RT_RENDER_FUSED_SHADE) != 0;
    sc->trace_unlit = (p.flags & RT_RENDER_TRACE_UNLIT) != 0;
    sc->sort_rays = (p.flags & RT_RENDER_SORT_RAYS) != 0;
    sc->split_shade = (p.flags & RT_RENDER_SPLIT_SHADE) != 0;
    sc->tile_kernel = (p.flags & RT_RENDER_TILE_KERNEL) != 0;
    sc->fast_accept = (p.flags & RT_RENDER_FAST_ACCEPT) != 0;
    // RT_RENDER_PEER_STORE: this frame's epoch and canvases (double-buffered by the parity of the epoch)
    uint32_t* peer_own = nullptr;
    PeerOut po;
    std::memset(&po, 0, sizeof po);
    if (peer_store) {
        const int n = sc->peer_n;
        if (n < 1 || p.world != n || (n > 1 && (p.rank < 0 || p.rank >= n)) || (p.flags & RT_RENDER_TILED_OUTPUT)) {
            set_error("RT_RENDER_PEER_STORE needs rt_scene_set_peers, world == its n_ranks and a row-major canvas (no RT_RENDER_TILED_OUTPUT)");
            return RT_ERR_INVALID_ARG;
        }
        p.flags |= RT_RENDER_CANVAS_DEVICE;
        const unsigned epoch = ++sc->peer_epoch;
        uint32_t* const* set = sc->peer_canv.data() + (size_t)(epoch & 1u) * n;
        peer_own = set[tm.rank];
        if (sc->peer_root >= 0) { po.pc.n = 1; po.pc.p[0] = set[sc->peer_root]; }
        else { po.pc.n = n; for (int r = 0; r < n; r++) po.pc.p[r] = set[r]; }
        if (!sc->d_push_ctr.p) {
            if (sc->d_push_ctr.reserve(16)) return RT_ERR_OUT_OF_MEMORY;
            RT_CUDA(cudaMemsetAsync(sc->d_push_ctr.p, 0, 16, st));
        }
        po.sg.mailbox_offset = sc->peer_pixels; po.sg.epoch = epoch; po.sg.done_blocks = sc->d_push_ctr.as<unsigned>();
        // under capture the kernels read the epoch from a pinned word the host rewrites before every replay
        *(volatile unsigned*)(sc->h_pinned + 49) = epoch;
        po.sg.epoch_ptr = sc->capturing ? (const unsigned*)(sc->h_pinned + 49) : nullptr;
        po.rank = tm.rank;
        canvas = peer_own;
    }
    const bool gather_sender = peer_store && sc->peer_root >= 0 && sc->peer_root != tm.rank && sc->peer_n > 1;
    bool ack_waited = false;
    {
        static const bool env_direct = [] { const char* e = std::getenv("RT_PEER_DIRECT"); return e && std::atoi(e); }();
        if (gather_sender && env_direct) {  // the frame kernels may store into the gathering rank themselves: wait in front of them
            k_wait_ack<<<1, 32, 0, st>>>(sc->peer_canv[tm.rank], sc->peer_pixels, sc->peer_root, po.sg.epoch, mailbox_timeout_cycles(),
                                         exchange_status(sc), po.sg.epoch_ptr);
            sc->kernel_launches++;
            ack_waited = true;
        }
    }
    bool built = false;
    RT_CUDA(rth::record(sc, sc->ev[2], st));
    if (!no_bvh) {
        if (int rc = ensure_bvh(sc, (p.flags & RT_RENDER_REBUILD_BVH) != 0, &built, (p.flags & RT_RENDER_REFIT_BVH) != 0)) return rc;
    } else {
        if (int rc = rth::update_instance_boxes(sc)) return rc;
    }
    long long out_pixels = tm.tiled_out ? slab_pixels(tm) : (long long)tm.W * tm.H;
    uint32_t* dcanvas = canvas;
    const bool host_canvas = !(p.flags & RT_RENDER_CANVAS_DEVICE);
    const bool host_bands = host_canvas && (p.flags & RT_RENDER_HOST_BANDS);
    if (host_bands && !(tm.tiled_out && tm.tiles_x == 1)) {
        set_error("RT_RENDER_HOST_BANDS needs RT_RENDER_TILED_OUTPUT and tile_w >= width (full-width bands)");
        return RT_ERR_INVALID_ARG;
    }
    // RT_RENDER_HOST_DIRECT: a host canvas in page-locked memory that the device can address (cudaHostAlloc /
    // cudaHostRegister, e.g. rt_host_canvas_open or a torch pinned tensor) is written by the kernels themselves, pixel
    // by pixel over PCIe while the frame renders: no staging canvas, no copy after the last kernel. Opt-in: measured on
    // B200 the scattered 4-byte stores to system memory stall the shading kernels (world16 at 4K: 3.36 ms against
    // 2.80 ms for the staged, chunked copy; N = 2: 1.87 against 1.77).
    bool direct = false;
    if (host_canvas && (p.flags & RT_RENDER_HOST_DIRECT)) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, canvas) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
            direct = true;
            dcanvas = (uint32_t*)at.devicePointer;
            if (host_bands) {  // every rank addresses the one row-major canvas; nothing of it is cleared here
                tm.tiled_out = 0;
                out_pixels = (long long)tm.W * tm.H;
            }
        } else {
            (void)cudaGetLastError();
        }
    }
    if (host_canvas && !direct) {
        if (sc->d_canvas.reserve(4 * (size_t)out_pixels)) return RT_ERR_OUT_OF_MEMORY;
        dcanvas = sc->d_canvas.as<uint32_t>();
    }
    const bool wavefront = !no_bvh && p.semantics == RT_SEMANTICS_GPU;
    // Host canvas: cut the frame into chunks of tile rows and copy chunk c while chunk c + 1 renders. Possible when a
    // chunk's pixels are whole rows of the host canvas: one rank with a row-major canvas, or full-width bands.
    int n_chunks = 1;
    const int units = tm.tiles_x > 0 ? tm.n_owned / tm.tiles_x : 0;  // tile rows (world 1) or bands (tiles_x == 1)
    if (wavefront && host_canvas && !direct && ((tm.world == 1 && !tm.tiled_out) || host_bands)) {
        static const int env_chunks = [] { const char* e = std::getenv("RT_HOST_CHUNKS"); return e ? std::atoi(e) : 0; }();
        // measured on B200 (world16 at 4K, 33 MB canvas): every extra chunk costs ~0.115 ms of kernel tails, the copy it
        // hides runs at ~57 GB/s; two chunks win from ~16 MiB of canvas per rank on
        // (not with secondary rays: every chunk would pay the cooperative bounce-wave launch and its grid barriers again —
        // world8_stress at 4K: 3.99 ms with two chunks against 2.87 + 0.58 without)
        const bool bounces = sc->any_secondary && sc->env.depth > 0;
        n_chunks = p.host_chunks > 0 ? p.host_chunks : env_chunks > 0 ? env_chunks
                   : (!bounces && 4 * (size_t)out_pixels >= ((size_t)16 << 20) ? 2 : 1);
        n_chunks = std::max(1, std::min(std::min(n_chunks, (int)RtScene::kMaxChunks), units));
    }
    // Concurrent lanes (RT_FRAME_LANES=n, default 1 = off): the owned tiles are cut into n contiguous ranges; each range runs
    // the wave-0 pipeline (k_tile<1> -> k_shadow_gen -> shadow kernel -> k_resolve) on its own stream with its own counter
    // block, events and slice of the queues. A traversal kernel ends with a few SMs finishing their longest work units while
    // the others idle (0.04-0.07 ms per kernel at any frame size, profiles/r2_summary.md section 5); with lanes the next
    // lane's kernel takes the SMs the tail leaves free. Same kernels, same arguments up to the tile range: same frame.
    // Only for the default pipeline of a device-resident frame without secondary rays, outside graph capture and stats
    // frames, and once the shadow-kernel tuner of this tile shape has settled (its timings need frames without overlap).
    int n_lanes = 1;
    int lane_bounds[RtScene::kMaxChunks + 1] = {0};
    {
        static const bool tile_env = [] { const char* e = std::getenv("RT_TILE_KERNEL"); return e && std::atoi(e); }();
        static const bool direct_env = [] { const char* e = std::getenv("RT_PEER_DIRECT"); return e && std::atoi(e); }();
        static const bool shadow_forced = std::getenv("RT_BEAM_SHADOW") != nullptr;
        static const int first_pct = [] { const char* e = std::getenv("RT_LANE_FIRST_PCT"); int v = e ? std::atoi(e) : 0; return (v > 0 && v < 100) ? v : 0; }();
        const int want = sc->lanes_req > 0 ? sc->lanes_req : frame_lanes_env();
        const bool bounces = sc->any_secondary && sc->env.depth > 0;
        if (want > 1 && wavefront && n_chunks == 1 && (p.flags & RT_RENDER_CANVAS_DEVICE) && !direct && !sc->capturing &&
            !(p.flags & RT_RENDER_GRAPH) && !want_stats && !bounces && !sc->any_refractive && !sc->fused_shade && !sc->sort_rays &&
            !sc->split_shade && !sc->tile_kernel && !tile_env && !direct_env && beam_primary_eligible(sc, TR_BVH_BOX) &&
            (shadow_forced || sc->tuners[sc->sb_cur].choice >= 0) && sc->n_lights > 0) {
            n_lanes = std::min(want, std::max(tm.n_owned / 64, 1));  // a lane is worth its launches from ~64 screen tiles on
            for (int c = 0; c <= n_lanes; c++) lane_bounds[c] = (int)((long long)tm.n_owned * c / n_lanes);
            if (n_lanes == 2 && first_pct) lane_bounds[1] = std::min(std::max((int)((long long)tm.n_owned * first_pct / 100), 1), tm.n_owned - 1);
        }
        if (n_lanes > 1) {
            if (int rc_l = ensure_lane_streams(sc, n_lanes)) return rc_l;
            // the queues of all lanes at once: a lane must not reallocate what another lane is using
            const size_t tot = (size_t)tm.n_work() * 32, nl = (size_t)std::max(sc->n_lights, 1);
            if (sc->d_hitq.reserve(sizeof(HitRec) * std::max<size_t>(tot, 1)) || sc->d_shadowq.reserve(sizeof(ShadowRec) * std::max<size_t>(tot, 1) * nl) ||
                sc->d_occl.reserve(std::max<size_t>(tot, 1) * nl) || sc->d_shq_count.reserve(sizeof(unsigned) * nl * (size_t)n_lanes))
                return RT_ERR_OUT_OF_MEMORY;
        }
    }
    const int n_blocks = std::max(n_chunks, n_lanes);  // counter blocks of this frame
    unsigned long long* ctr = sc->d_counters.as<unsigned long long>();
    RT_CUDA(cudaMemsetAsync(ctr, 0, sizeof(unsigned long long) * C_N * n_blocks, st));
    // row-major canvas shared with other ranks' tiles: clear them. A tile slab needs no clearing: every
    // in-canvas pixel of an owned tile is written, and the slab padding beyond the canvas edge is never read.
    // (not a peer canvas either: the other ranks are storing their tiles into it)
    if (!tm.tiled_out && tm.world > 1 && !(direct && host_bands) && !peer_store) RT_CUDA(cudaMemsetAsync(dcanvas, 0, 4 * (size_t)out_pixels, st));
    RT_CUDA(rth::record(sc, sc->ev[3], st));
    SceneView v = rth::make_view(sc, !no_bvh);
    int waves = 1, rc = RT_OK;
    bool canvas_copied = false, peer_direct = false;
    if (no_bvh) rc = run_mega<TR_BRUTE>(sc, v, tm, dcanvas, p.semantics);
    else if (p.semantics == RT_SEMANTICS_CPU) rc = run_mega<TR_BVH_BOX>(sc, v, tm, dcanvas, p.semantics);
    else if (n_lanes > 1) {
        // RT_LANE_ORDER=1: the last lane (the bottom of the image, where the long work units are) is enqueued first
        static const bool bottom_first = [] { const char* e = std::getenv("RT_LANE_ORDER"); return e && std::atoi(e); }();
        RT_CUDA(cudaEventRecord(sc->lane_fork, st));
        for (int i = 0; i < n_lanes && !rc; i++) {
            const int c = bottom_first ? n_lanes - 1 - i : i;
            cudaStream_t ls = c == 0 ? st : sc->lane_stream[c - 1];
            if (c > 0) RT_CUDA(cudaStreamWaitEvent(ls, sc->lane_fork, 0));
            TileMap tc = tm;
            tc.k_begin = lane_bounds[c]; tc.k_count = lane_bounds[c + 1] - lane_bounds[c];
            sc->cur_stream = ls; sc->cur_lane = c;
            sc->cur_off = (long long)tc.k_begin * tm.sub_per_tile * 32;
            int w = 1;
            if (v.fast_accept) rc = run_wavefront<TR_BVH_BOX_FAST, false>(sc, v, tc, dcanvas, &w, ctr + (size_t)C_N * c, sc->cev[c], nullptr);
            else rc = run_wavefront<TR_BVH_BOX, false>(sc, v, tc, dcanvas, &w, ctr + (size_t)C_N * c, sc->cev[c], nullptr);
            if (!rc && c > 0) {
                cudaError_t e_j = cudaEventRecord(sc->lane_join[c - 1], ls);
                if (e_j != cudaSuccess) rc = rth::cuda_fail(e_j, "cudaEventRecord(lane_join)", __FILE__, __LINE__);
            }
        }
        sc->cur_stream = nullptr; sc->cur_off = 0; sc->cur_lane = 0;
        for (int c = 1; c < n_lanes && !rc; c++) RT_CUDA(cudaStreamWaitEvent(st, sc->lane_join[c - 1], 0));
    } else {
        // chunk sizes fall geometrically (1, 1/4, 1/16, ...): only the last chunk's copy is not hidden behind rendering
        // (world16 at 4K, two chunks: e2e 2.496 ms at ratio 1/2, 2.457 at 1/4, 2.461 at 0.18).
        // Every chunk gets at least one unit (n_chunks <= units).
        int bounds[RtScene::kMaxChunks + 1];
        bounds[0] = 0;
        // ratio between consecutive chunk sizes (RT_CHUNK_RATIO overrides it for experiments)
        static const double ratio = [] { const char* e = std::getenv("RT_CHUNK_RATIO"); double r = e ? std::atof(e) : 0.0; return (r > 0.0 && r < 1.0) ? r : 0.25; }();
        const double tot = (1.0 - std::pow(ratio, n_chunks)) / (1.0 - ratio);
        for (int c = 1; c < n_chunks; c++) {
            int b = (int)(units * (((1.0 - std::pow(ratio, c)) / (1.0 - ratio)) / tot));
            bounds[c] = std::min(std::max(b, bounds[c - 1] + 1), units - (n_chunks - c));
        }
        bounds[n_chunks] = units;
        for (int c = 0; c < n_chunks && !rc; c++) {
            TileMap tc = tm;
            const int u0 = bounds[c], u1 = bounds[c + 1];
            if (n_chunks > 1) { tc.k_begin = u0 * tm.tiles_x; tc.k_count = (u1 - u0) * tm.tiles_x; }
            int w = 1;
            // peer frames without secondary rays: k_tile stores into the peers and signals them itself
            const bool secondary = v.any_secondary && v.depth > 0;
            // Opt-in (RT_PEER_DIRECT=1): measured on 8 B200 the 4-byte pixel stores of a warp tile (four 32-byte runs) over
            // NVLink stall the tile kernels — 0.333 ms for 1/8 of a world16 4K frame against 0.26 ms with local stores plus
            // the 16-byte-vector forward kernel below (profiles/r2_summary.md) — the same finding as RT_RENDER_HOST_DIRECT.
            static const bool env_direct = [] { const char* e = std::getenv("RT_PEER_DIRECT"); return e && std::atoi(e); }();
            peer_direct = env_direct && peer_store && !secondary && tile_kernel_eligible(sc, v, TR_BVH_BOX);
            const PeerOut* pp = peer_direct ? &po : nullptr;
            if (v.fast_accept) {
                if (want_stats) rc = run_wavefront<TR_BVH_BOX_FAST, true>(sc, v, tc, dcanvas, &w, ctr + (size_t)C_N * c, sc->cev[c], pp);
                else rc = run_wavefront<TR_BVH_BOX_FAST, false>(sc, v, tc, dcanvas, &w, ctr + (size_t)C_N * c, sc->cev[c], pp);
            } else if (want_stats) rc = run_wavefront<TR_BVH_BOX, true>(sc, v, tc, dcanvas, &w, ctr + (size_t)C_N * c, sc->cev[c], pp);
            else rc = run_wavefront<TR_BVH_BOX, false>(sc, v, tc, dcanvas, &w, ctr + (size_t)C_N * c, sc->cev[c], pp);
            waves = (w < 0 || waves < 0) ? -1 : std::max(waves, w);
            if (rc || n_chunks == 1) continue;
            // this chunk's rows go to the host while the next chunk renders
            RT_CUDA(cudaEventRecord(sc->chunk_done[c], st));
            RT_CUDA(cudaStreamWaitEvent(sc->copy_stream, sc->chunk_done[c], 0));
            if (int rc2 = copy_rows_to_host(sc, tm, tc.k_begin, tc.k_count, dcanvas, canvas, host_bands, sc->copy_stream)) return rc2;
        }
        if (!rc && n_chunks > 1) {
            RT_CUDA(cudaEventRecord(sc->copies_done, sc->copy_stream));
            RT_CUDA(cudaStreamWaitEvent(st, sc->copies_done, 0));
            canvas_copied = true;
        }
    }
    if (rc) return rc;
    if (peer_store) {
        if (!peer_direct) {  // the frame is in this rank's own canvas: forward the owned tiles and signal
            if (tm.tile_w % 4) { set_error("tile_w must be a multiple of 4"); return RT_ERR_INVALID_ARG; }
            const long long n_quads = (long long)tm.n_owned * tm.tile_h * (tm.tile_w / 4);
            const unsigned grid = (unsigned)std::max<long long>((n_quads + kThreads - 1) / kThreads, 1);
            if (gather_sender && !ack_waited) {  // sender of a gather: not more than one frame ahead of the gathering rank
                k_wait_ack<<<1, 32, 0, st>>>(sc->peer_canv[tm.rank], sc->peer_pixels, sc->peer_root, po.sg.epoch, mailbox_timeout_cycles(),
                                             exchange_status(sc), po.sg.epoch_ptr);
                sc->kernel_launches++;
            }
            k_forward_tiles<<<grid, kThreads, 0, st>>>(tm, peer_own, po.pc, po.sg);
            sc->kernel_launches++;
        }
        if (gather_sender) {  // rt_frame_finish looks at the time-out word of the acknowledgement wait
            RT_CUDA(rth::record(sc, sc->ev_exchange, st));
            sc->exchange_pending = true;
        }
        if (sc->peer_root < 0 || sc->peer_root == tm.rank) {  // this rank receives the frame: wait until every rank has signalled
            PeerCanvases ack;
            std::memset(&ack, 0, sizeof ack);
            if (sc->peer_root >= 0 && sc->peer_n > 1) { ack.n = sc->peer_n; for (int r = 0; r < sc->peer_n; r++) ack.p[r] = sc->peer_canv[r]; }  // parity-0 canvases hold the acknowledgement words
            k_wait_mailbox<<<1, 32, 0, st>>>(peer_own, sc->peer_pixels, sc->peer_n, po.sg.epoch, mailbox_timeout_cycles(), exchange_status(sc), po.sg.epoch_ptr,
                                             ack, std::max(sc->peer_root, 0));
            sc->kernel_launches++;
            RT_CUDA(rth::record(sc, sc->ev_exchange, st));
            sc->exchange_pending = true;
        }
        RT_CUDA(cudaGetLastError());
    }
    RT_CUDA(rth::record(sc, sc->ev[4], st));
    unsigned long long* host_ctr = sc->h_pinned + 64;
    RT_CUDA(cudaMemcpyAsync(host_ctr, ctr, sizeof(unsigned long long) * C_N * n_blocks, cudaMemcpyDeviceToHost, st));
    if (host_canvas && !canvas_copied && !direct) {
        if (host_bands) {
            if (int rc2 = copy_rows_to_host(sc, tm, 0, tm.n_owned, dcanvas, canvas, true, st)) return rc2;
        } else {
            RT_CUDA(cudaMemcpyAsync(canvas, dcanvas, 4 * (size_t)out_pixels, cudaMemcpyDeviceToHost, st));
        }
    }
    sc->pend_chunks = n_blocks;
    sc->pend_lanes = n_lanes;
    sc->pend_direct = direct;
    RT_CUDA(rth::record(sc, sc->ev[5], st));
    sc->pending = true;
    sc->pend_stats = want_stats;
    sc->pend_built = built; sc->pend_wavefront = wavefront; sc->pend_waves = waves; sc->pend_launches0 = launches0;
    return RT_OK;
}

// ---- RT_RENDER_GRAPH: the frame as one CUDA graph
// Everything a captured frame bakes into kernel parameters or launch shapes. Two frames with equal keys enqueue identical work.
static void graph_key(const RtScene* sc, const RtRenderParams& p, const uint32_t* canvas, std::vector<unsigned char>& key) {
    key.clear();
    auto put = [&](const void* d, size_t n) { const unsigned char* b = (const unsigned char*)d; key.insert(key.end(), b, b + n); };
    put(&p, sizeof p);
    put(&canvas, sizeof canvas);
    put(&sc->width, sizeof sc->width); put(&sc->height, sizeof sc->height);
    put(&sc->camera, sizeof sc->camera); put(&sc->env, sizeof sc->env);
    put(&sc->inst_gen, sizeof sc->inst_gen); put(&rth::g_alloc_gen, sizeof rth::g_alloc_gen);
    const int st[8] = {sc->boxes_valid, sc->bvh.valid, sc->bvh.topology_valid, sc->tuners[sc->sb_cur].choice, sc->n_instances, sc->peer_n, sc->peer_root,
                       (p.flags & RT_RENDER_PEER_STORE) ? (int)((sc->peer_epoch + 1u) & 1u) : 0};
    put(st, sizeof st);
    put(&sc->stream, sizeof sc->stream);
    const size_t n_peer = sc->peer_canv.size();
    if (n_peer) put(sc->peer_canv.data(), n_peer * sizeof(uint32_t*));
}
static rth::FrameState save_state(const RtScene* sc, int launches) {
    rth::FrameState f;
    f.pend_built = sc->pend_built; f.pend_wavefront = sc->pend_wavefront; f.pend_stats = sc->pend_stats; f.pend_direct = sc->pend_direct;
    f.split_shade_used = sc->split_shade_used; f.tile_kernel_used = sc->tile_kernel_used; f.exchange_pending = sc->exchange_pending;
    f.pend_waves = sc->pend_waves; f.pend_chunks = sc->pend_chunks; f.wave0_form = sc->wave0_form; f.sb_used = sc->sb_used; f.launches = launches;
    return f;
}
static void restore_state(RtScene* sc, const rth::FrameState& f) {
    sc->pend_built = f.pend_built; sc->pend_wavefront = f.pend_wavefront; sc->pend_stats = f.pend_stats; sc->pend_direct = f.pend_direct;
    sc->split_shade_used = f.split_shade_used; sc->tile_kernel_used = f.tile_kernel_used; sc->exchange_pending = f.exchange_pending;
    sc->pend_waves = f.pend_waves; sc->pend_chunks = f.pend_chunks; sc->wave0_form = f.wave0_form; sc->sb_used = f.sb_used;
    sc->pend_lanes = 1;  // captured frames never use lanes
}

int rt_render(RtScene* sc, const RtRenderParams* params, uint32_t* canvas, RtFrameStats* stats) {
    if (!sc) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    RtRenderParams p;
    std::memset(&p, 0, sizeof p);
    if (params) p = *params;
    if (sc->pending) {  // an asynchronous frame was never finished: its counters are about to be reused
        if (int rc = rt_frame_finish(sc, nullptr)) return rc;
    }
    {   // the shadow-kernel tuner of this frame's tile shape
        const int shape[5] = {p.tile_w, p.tile_h, p.world > 1 ? p.world : 1, sc->width, sc->height};
        int found = -1;
        for (int i = 0; i < 6 && found < 0; i++) if (sc->tuners[i].used && std::memcmp(sc->tuners[i].shape, shape, sizeof shape) == 0) found = i;
        if (found < 0) {
            found = sc->sb_next;
            sc->sb_next = (sc->sb_next + 1) % 6;
            sc->tuners[found] = RtScene::SbTuner();
            std::memcpy(sc->tuners[found].shape, shape, sizeof shape);
            sc->tuners[found].used = true;
        }
        sc->sb_cur = found;
    }
    // A frame can be a graph when nothing in it needs the host in between: BVH traversal with the CUDA bounce semantics
    // (wavefront kernels), canvas on the device or in the peers, and the wave-0 shadow kernel already chosen.
    const bool peer_store = (p.flags & RT_RENDER_PEER_STORE) != 0;
    static const int graph_env = [] { const char* e = std::getenv("RT_GRAPH"); return e ? (std::atoi(e) ? 1 : 0) : -1; }();
    const bool want_graph = graph_env >= 0 ? graph_env == 1 : (p.flags & RT_RENDER_GRAPH) != 0;
    const bool graphable = want_graph && !(p.flags & RT_RENDER_NO_BVH) && p.semantics == RT_SEMANTICS_GPU &&
                           ((p.flags & RT_RENDER_CANVAS_DEVICE) || peer_store) && (canvas || peer_store);
    int rc = RT_OK, pend_graph = 0;
    bool done = false;
    if (graphable) {
        std::vector<unsigned char> key;
        graph_key(sc, p, canvas, key);
        const int launches0 = sc->kernel_launches;
        rth::FrameGraph* hit = nullptr;
        for (rth::FrameGraph& g : sc->graphs) if (g.exec && g.key == key) hit = &g;
        if (hit) {
            // replay: the host-side effects of an enqueue, then one launch
            if (peer_store) { sc->peer_epoch++; *(volatile unsigned*)(sc->h_pinned + 49) = sc->peer_epoch; }
            restore_state(sc, hit->state);
            sc->boxes_valid = true;
            if (hit->state.pend_built) { sc->bvh.valid = true; sc->bvh.topology_valid = true; }
            RT_CUDA(cudaGraphLaunch(hit->exec, sc->stream));
            sc->kernel_launches += hit->state.launches;
            sc->pend_launches0 = launches0;
            sc->pending = true;
            hit->last_used = ++sc->graph_clock;
            sc->graph_replays++;
            pend_graph = 2;
            done = true;
        } else if ((key == sc->last_key || key == sc->prev_key) && key != sc->bad_key && (sc->sb_used < 0 || sc->tuners[sc->sb_cur].choice >= 0)) {
            // second appearance of this key (peer frames alternate between the keys of the two canvas parities, hence the last two
            // frames): every buffer has its size, capture the frame
            const unsigned epoch0 = sc->peer_epoch;
            const bool boxes0 = sc->boxes_valid, valid0 = sc->bvh.valid, topo0 = sc->bvh.topology_valid;
            cudaGraph_t graph = nullptr;
            cudaError_t e = cudaStreamBeginCapture(sc->stream, cudaStreamCaptureModeRelaxed);
            if (e == cudaSuccess) {
                sc->capturing = true;
                const int rc_cap = render_enqueue(sc, &p, canvas);
                sc->capturing = false;
                e = cudaStreamEndCapture(sc->stream, &graph);
                cudaGraphExec_t exec = nullptr;
                if (rc_cap == RT_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&exec, graph, 0);
                if (graph) cudaGraphDestroy(graph);
                if (rc_cap == RT_OK && e == cudaSuccess && exec) {
                    rth::FrameGraph* slot = nullptr;  // an empty entry, else the least recently used one
                    for (rth::FrameGraph& g : sc->graphs) if (!g.exec) { slot = &g; break; }
                    if (!slot) { slot = &sc->graphs[0]; for (rth::FrameGraph& g : sc->graphs) if (g.last_used < slot->last_used) slot = &g; }
                    if (slot->exec) cudaGraphExecDestroy(slot->exec);
                    slot->exec = exec; slot->key = key; slot->state = save_state(sc, sc->kernel_launches - launches0); slot->last_used = ++sc->graph_clock;
                    RT_CUDA(cudaGraphLaunch(exec, sc->stream));
                    sc->graph_captures++;
                    pend_graph = 1;
                    done = true;
                } else {
                    // not capturable (an operation in this frame needs the host): plain launches for this key from now on
                    if (exec) cudaGraphExecDestroy(exec);
                    (void)cudaGetLastError();
                    // nothing of the captured frame ran on the device: take back what its enqueue claimed on the host
                    sc->pending = false;
                    sc->exchange_pending = false;
                    sc->peer_epoch = epoch0;
                    sc->kernel_launches = launches0;
                    sc->boxes_valid = boxes0; sc->bvh.valid = valid0; sc->bvh.topology_valid = topo0;
                    sc->bad_key = key;
                }
            } else {
                (void)cudaGetLastError();
                sc->bad_key = key;
            }
        }
        sc->prev_key.swap(sc->last_key);
        sc->last_key.swap(key);
    }
    if (!done) rc = render_enqueue(sc, &p, canvas);
    if (rc) return rc;
    sc->pend_graph = pend_graph;
    if (p.flags & RT_RENDER_ASYNC) return RT_OK;  // rt_frame_finish() completes the frame
    return rt_frame_finish(sc, stats);
}

int rt_frame_finish(RtScene* sc, RtFrameStats* stats) {
    if (!sc) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    if (!sc->pending) { set_error("no frame pending"); return RT_ERR_INVALID_ARG; }
    sc->pending = false;
    const bool built = sc->pend_built, wavefront = sc->pend_wavefront;
    const int waves = sc->pend_waves, launches0 = sc->pend_launches0;
    const int n_chunks = sc->pend_chunks;
    unsigned long long host_ctr[C_N];  // counters summed over the frame's chunks
    unsigned long long max_waves = 0;
    for (int i = 0; i < C_N; i++) host_ctr[i] = 0;
    RtFrameStats fs;
    std::memset(&fs, 0, sizeof fs);
    RT_CUDA(cudaEventSynchronize(sc->ev[5]));
    // an exchange enqueued behind the frame on the same stream (rt_push_tiles_signal with RT_RENDER_ASYNC) reports its
    // time-out here once the stream has drained
    if (sc->exchange_pending) {
        sc->exchange_pending = false;
        RT_CUDA(cudaEventSynchronize(sc->ev_exchange));
        if (int rc_x = check_exchange(sc)) return rc_x;
    }
    for (int c = 0; c < n_chunks; c++) {
        const unsigned long long* hc = sc->h_pinned + 64 + (size_t)C_N * c;
        for (int i = 0; i < C_N; i++) host_ctr[i] += hc[i];
        max_waves = std::max(max_waves, hc[C_WAVES]);
    }
    host_ctr[C_WAVES] = max_waves;
    if (built) {
        RtBuildStats bs;
        if (int rc2 = rth::finish_build(sc, &bs)) return rc2;
        fs.ms_build = bs.ms_build;
    }
    fs.rays_primary = host_ctr[C_RAYS_PRIMARY];
    fs.rays_culled = host_ctr[C_RAYS_CULLED];
    fs.rays_shadow = host_ctr[C_RAYS_SHADOW] + fs.rays_culled;
    fs.rays_secondary = host_ctr[C_RAYS_SECONDARY];
    fs.rays_total = fs.rays_primary + fs.rays_shadow + fs.rays_secondary;
    fs.nodes_visited = host_ctr[C_NODES];
    fs.prims_prefiltered = host_ctr[C_PRIMS];
    fs.tris_tested = host_ctr[C_EXACT];
    RT_CUDA(cudaEventElapsedTime(&fs.ms_trace, sc->ev[3], sc->ev[4]));
    RT_CUDA(cudaEventElapsedTime(&fs.ms_total, sc->ev[2], sc->ev[5]));
    if (wavefront) {
        for (int c = 0; c < n_chunks; c++) {  // per-kernel times: summed over the chunks
            float ms = 0.0f;
            RT_CUDA(cudaEventElapsedTime(&ms, sc->cev[c][0], sc->cev[c][1])); fs.ms_primary += ms;
            RT_CUDA(cudaEventElapsedTime(&ms, sc->cev[c][1], sc->cev[c][4])); fs.ms_shade += ms;
            if (sc->split_shade_used) { RT_CUDA(cudaEventElapsedTime(&ms, sc->cev[c][2], sc->cev[c][3])); fs.ms_shadow += ms; }
            RT_CUDA(cudaEventElapsedTime(&ms, sc->cev[c][4], sc->cev[c][5])); fs.ms_bounce += ms;
        }
        fs.n_hits = (int32_t)(host_ctr[C_HITQ] >> 32);  // high word = hits, low word = queue slots
    }
    RtScene::SbTuner& tuner = sc->tuners[sc->sb_cur];
    if (wavefront && sc->split_shade_used && n_chunks == 1 && tuner.choice < 0 && sc->sb_used >= 0 && !sc->pend_stats) {
        // auto-selection of the wave-0 shadow kernel (run_wavefront, trace_shadows): best of two frames each, per tile shape
        float& best = tuner.ms[sc->sb_used];
        if (best <= 0.0f || fs.ms_shadow < best) best = fs.ms_shadow;
        if (++tuner.frames >= 6) {
            tuner.choice = 1;
            if (tuner.ms[2] < tuner.ms[tuner.choice]) tuner.choice = 2;
            if (tuner.ms[0] < tuner.ms[tuner.choice]) tuner.choice = 0;
        }
    }
    for (int k = 0; k < 4; k++) fs.ms_wave_phase[k] = (float)host_ctr[C_PHASE_NS + k] * 1e-6f;
    fs.waves = waves >= 0 ? waves : 1 + (int)host_ctr[C_WAVES];
    if (fs.waves > 4096) { set_error("bounce waves did not terminate"); return RT_ERR_LIMIT; }
    fs.kernel_launches = sc->kernel_launches - launches0;
    fs.host_chunks = sc->pend_direct ? 0 : sc->pend_lanes > 1 ? 1 : n_chunks;
    fs.tile_kernel = sc->wave0_form;
    fs.graph = sc->pend_graph;
    fs.node_fetches = host_ctr[C_FETCHES];
    if (stats) *stats = fs;
    return RT_OK;
}

int rt_update_scene(RtScene* sc, int32_t kernel_dim, int32_t optimize, uint32_t* host_canvas) {
    (void)kernel_dim;  // the reference's block edge (-d); grid shapes here are fixed by the SM count
    RtRenderParams p;
    std::memset(&p, 0, sizeof p);
    p.flags = optimize ? RT_RENDER_REBUILD_BVH : RT_RENDER_NO_BVH;  // the reference rebuilds every frame (src/raytracer.cu:104-107)
    p.semantics = RT_SEMANTICS_GPU;
    return rt_render(sc, &p, host_canvas, nullptr);
}

static int tracer_mode(RtScene* sc, int use_bvh, int* mode) {
    switch (use_bvh) {
        case 1: *mode = TR_BVH_BOX; break;
        case 0: *mode = TR_BRUTE; break;
        case 2: *mode = TR_BVH_NOBOX; break;
        case 3: *mode = TR_BRUTE_BOX; break;
        default: set_error("use_bvh must be 0 (brute force), 1 (BVH8 + leaf-box rule), 2 (BVH8 only), 3 (brute force + leaf-box rule)"); return RT_ERR_INVALID_ARG;
    }
    if (*mode == TR_BVH_BOX || *mode == TR_BVH_NOBOX) return ensure_bvh(sc, false, nullptr);
    return rth::update_instance_boxes(sc);
}

int rt_trace_primary(RtScene* sc, int32_t use_bvh, int32_t* instance, int32_t* triangle, float* t) {
    if (!sc || !instance || !triangle || !t) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    int mode;
    if (int rc = tracer_mode(sc, use_bvh, &mode)) return rc;
    size_t n = (size_t)sc->width * sc->height;
    if (sc->d_ids.reserve(12 * n)) return RT_ERR_OUT_OF_MEMORY;
    int* di = sc->d_ids.as<int>();
    int* dt = di + n;
    float* dtt = (float*)(di + 2 * n);
    TileMap tm;
    if (int rc = make_tilemap(sc, nullptr, tm)) return rc;
    unsigned long long* ctr = sc->d_counters.as<unsigned long long>();
    cudaStream_t st = sc->stream;
    RT_CUDA(cudaMemsetAsync(ctr, 0, sizeof(unsigned long long) * C_N, st));
    SceneView v = rth::make_view(sc, true);
    int g = sc->sm_count * 4;
    switch (mode) {
        case TR_BVH_BOX:
            if (v.fast_accept) k_primary<TR_BVH_BOX_FAST, false, true><<<g, kThreads, 0, st>>>(v, tm, ctr, nullptr, nullptr, di, dt, dtt, 0);
            else k_primary<TR_BVH_BOX, false, true><<<g, kThreads, 0, st>>>(v, tm, ctr, nullptr, nullptr, di, dt, dtt, 0);
            break;
        case TR_BVH_NOBOX: k_primary<TR_BVH_NOBOX, false, true><<<g, kThreads, 0, st>>>(v, tm, ctr, nullptr, nullptr, di, dt, dtt, 0); break;
        case TR_BRUTE: k_primary<TR_BRUTE, false, true><<<g, kThreads, 0, st>>>(v, tm, ctr, nullptr, nullptr, di, dt, dtt, 0); break;
        default: k_primary<TR_BRUTE_BOX, false, true><<<g, kThreads, 0, st>>>(v, tm, ctr, nullptr, nullptr, di, dt, dtt, 0); break;
    }
    sc->kernel_launches++;
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(instance, di, 4 * n, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(triangle, dt, 4 * n, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(t, dtt, 4 * n, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}

int rt_cast_rays(RtScene* sc, int32_t use_bvh, const float* origins, const float* dirs, int32_t n, int32_t* instance,
                 int32_t* triangle, float* t) {
    if (!sc || n < 0 || (n && (!origins || !dirs || !instance || !triangle || !t))) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    if (n == 0) return RT_OK;
    int mode;
    if (int rc = tracer_mode(sc, use_bvh, &mode)) return rc;
    size_t nn = (size_t)n;
    if (sc->d_ids.reserve(12 * nn + 24 * nn)) return RT_ERR_OUT_OF_MEMORY;
    int* di = sc->d_ids.as<int>();
    int* dt = di + nn;
    float* dtt = (float*)(di + 2 * nn);
    float* d_o = dtt + nn;
    float* d_d = d_o + 3 * nn;
    cudaStream_t st = sc->stream;
    RT_CUDA(cudaMemcpyAsync(d_o, origins, 12 * nn, cudaMemcpyHostToDevice, st));
    RT_CUDA(cudaMemcpyAsync(d_d, dirs, 12 * nn, cudaMemcpyHostToDevice, st));
    SceneView v = rth::make_view(sc, true);
    int g = (n + kThreads - 1) / kThreads;
    switch (mode) {
        case TR_BVH_BOX: k_cast_rays<TR_BVH_BOX><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
        case TR_BVH_NOBOX: k_cast_rays<TR_BVH_NOBOX><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
        case TR_BRUTE: k_cast_rays<TR_BRUTE><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
        default: k_cast_rays<TR_BRUTE_BOX><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
    }
    sc->kernel_launches++;
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(instance, di, 4 * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(triangle, dt, 4 * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(t, dtt, 4 * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}

int rt_debug_cast(RtScene* sc, int32_t x, int32_t y, int32_t semantics, RtHit* out) {
    if (!sc || !out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    if (x < 0 || y < 0 || x >= sc->width || y >= sc->height) { set_error("pixel outside the canvas"); return RT_ERR_INVALID_ARG; }
    if (int rc = ensure_bvh(sc, false, nullptr)) return rc;
    if (sc->d_ids.reserve(sizeof(RtHit))) return RT_ERR_OUT_OF_MEMORY;
    SceneView v = rth::make_view(sc, true);
    cudaStream_t st = sc->stream;
    if (semantics == RT_SEMANTICS_CPU) {
        RT_CUDA(cudaDeviceSetLimit(cudaLimitStackSize, 32 * 1024));
        k_debug<TR_BVH_BOX, RT_SEMANTICS_CPU><<<1, 1, 0, st>>>(v, x, y, sc->d_ids.as<RtHit>());
        RT_CUDA(cudaStreamSynchronize(st));
        RT_CUDA(cudaDeviceSetLimit(cudaLimitStackSize, 2048));
    } else {
        k_debug<TR_BVH_BOX, RT_SEMANTICS_GPU><<<1, 1, 0, st>>>(v, x, y, sc->d_ids.as<RtHit>());
    }
    sc->kernel_launches++;
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(out, sc->d_ids.p, sizeof(RtHit), cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}

int rt_debug_get(RtScene* sc, int32_t what, int64_t* out) {
    if (!sc || !out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
    switch (what) {
        case RT_DEBUG_PREFIX_UPLOADS: *out = sc->bvh.prefix_uploads; break;
        case RT_DEBUG_BVH8_LEVELS: *out = sc->bvh.n_levels8; break;
        case RT_DEBUG_BUILD_PATH: *out = sc->bvh.build_path; break;
        case RT_DEBUG_BUILD_LAUNCHES: *out = sc->bvh.build_launches; break;
        case RT_DEBUG_GRAPH_REPLAYS: *out = sc->graph_replays; break;
        case RT_DEBUG_GRAPH_CAPTURES: *out = sc->graph_captures; break;
        case RT_DEBUG_FRAME_LANES: *out = sc->pend_lanes; break;
        case RT_DEBUG_BUILD_CLUSTER: *out = sc->bvh.build_path == 2 ? sc->bvh.cluster_size : 0; break;
        default:
            // RT_DEBUG_BUILD_PHASE_NS + i, i in 0..6: nanoseconds CTA 0 of the last cluster build spent in phase i (flatten +
            // bounds, Morton, sort, Karras, chunk-local boxes, spanning boxes, BVH8 collapse); needs RT_BUILD_TIMERS=1
            if (what >= RT_DEBUG_BUILD_LEVEL_NS && what < RT_DEBUG_BUILD_LEVEL_NS + 32) {  // raw stamp i of the collapse levels
                unsigned long long v = 0;
                if (!sc->bvh.scene_bounds.p || sc->bvh.scene_bounds.cap < 512) { set_error("no cluster build yet"); return RT_ERR_NO_BVH; }
                RT_CUDA(cudaMemcpy(&v, (const char*)sc->bvh.scene_bounds.p + 256 + 8 * (what - RT_DEBUG_BUILD_LEVEL_NS), 8, cudaMemcpyDeviceToHost));
                *out = (int64_t)v;
                break;
            }
            if (what >= RT_DEBUG_BUILD_PHASE_NS && what < RT_DEBUG_BUILD_PHASE_NS + 7) {
                const unsigned long long* t = sc->h_pinned + 50;
                *out = (int64_t)(t[what - RT_DEBUG_BUILD_PHASE_NS + 1] - t[what - RT_DEBUG_BUILD_PHASE_NS]);
                break;
            }
            set_error("unknown debug counter"); return RT_ERR_INVALID_ARG;
    }
    return RT_OK;
}

int rt_measure_l2_bandwidth(int64_t bytes, int32_t iters, float* gbps) {
    if (!gbps || bytes < 4096 || iters < 1) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    if (g_device < 0) { if (int rc = rt_init(0)) return rc; }
    void* buf = nullptr;
    unsigned* sink = nullptr;
    RT_CUDA(cudaMalloc(&buf, (size_t)bytes));
    RT_CUDA(cudaMalloc((void**)&sink, 4));
    RT_CUDA(cudaMemset(buf, 1, (size_t)bytes));
    cudaEvent_t e0, e1;
    RT_CUDA(cudaEventCreate(&e0));
    RT_CUDA(cudaEventCreate(&e1));
    long long n16 = bytes / 16;
    int grid = g_sm_count * 8;
    k_l2_read<<<grid, kThreads>>>((const uint4*)buf, n16, 2, sink);  // warm the cache
    RT_CUDA(cudaEventRecord(e0));
    k_l2_read<<<grid, kThreads>>>((const uint4*)buf, n16, iters, sink);
    RT_CUDA(cudaEventRecord(e1));
    RT_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    RT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    *gbps = (float)((double)n16 * 16.0 * iters / (ms * 1e-3) / 1e9);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    cudaFree(sink);
    return RT_OK;
}

}  // extern "C"

// ================================================================ rt_cast_rays without any allocation
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). rt_cast_rays stages the rays and their
// results in RtScene::d_ids, a grow-only cudaMalloc buffer; this variant runs the same kernel (k_cast_rays) over a
// file-scope device array instead: 36 bytes per ray (origin 12, direction 12, instance 4, triangle 4, t 4) for at most
// RT_CAST_STATIC_MAX_RAYS rays = 2.25 MiB, held by every GPU that loads the library until the process exits.
//   * In host code `g_cast_static` is only a host-side stand-in for the symbol, not a device address: its device address is
//     looked up once per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernel gets.
//   * Every call — any scene, any stream, any host thread — shares the one array of its GPU. It keeps the previous call's
//     contents; nothing in it has to start at zero, because every word a call reads was written by that same call (the
//     inputs are copied into [0, n), the kernel stores all three results for every i < n). Calls that overlap on the GPU
//     race; the entry point synchronises its stream before it returns and serialises its callers on the host.
//   * No cudaMalloc / cudaFree / allocating library call: a BVH has to exist already (RT_ERR_NO_BVH otherwise).
#include <mutex>

static __device__ __align__(256) unsigned char g_cast_static[(size_t)RT_CAST_STATIC_MAX_RAYS * 36];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_cast_static_mutex;
void* g_cast_static_ptr[kMaxDevices] = {};  // device address of g_cast_static per device ordinal, null = not looked up yet

// Set-up, once per device, with that device current. Called with g_cast_static_mutex held.
int cast_static_address(unsigned char** out) {
    int dev = -1;
    RT_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { set_error("device ordinal out of range for the static scratch table"); return RT_ERR_LIMIT; }
    if (!g_cast_static_ptr[dev]) {
        void* p = nullptr;
        RT_CUDA(cudaGetSymbolAddress(&p, g_cast_static));
        if (!p) { set_error("cudaGetSymbolAddress returned a null address for the static scratch array"); return RT_ERR_CUDA; }
        g_cast_static_ptr[dev] = p;
    }
    *out = (unsigned char*)g_cast_static_ptr[dev];
    return RT_OK;
}
}  // namespace

extern "C" int rt_cast_rays_static(RtScene* sc, int32_t use_bvh, const float* origins, const float* dirs, int32_t n,
                                   int32_t* instance, int32_t* triangle, float* t) {
    if (!sc || n < 0 || (n && (!origins || !dirs || !instance || !triangle || !t))) { set_error("invalid argument"); return RT_ERR_INVALID_ARG; }
    if (n > RT_CAST_STATIC_MAX_RAYS) {
        set_error("rt_cast_rays_static: n = " + std::to_string(n) + " rays exceeds RT_CAST_STATIC_MAX_RAYS = " +
                  std::to_string((int)RT_CAST_STATIC_MAX_RAYS) + " (the static scratch array); split the batch or use rt_cast_rays");
        return RT_ERR_LIMIT;
    }
    int mode;
    switch (use_bvh) {
        case 1: mode = TR_BVH_BOX; break;
        case 0: mode = TR_BRUTE; break;
        case 2: mode = TR_BVH_NOBOX; break;
        case 3: mode = TR_BRUTE_BOX; break;
        default: set_error("use_bvh must be 0, 1, 2 or 3 (see rt_cast_rays)"); return RT_ERR_INVALID_ARG;
    }
    if (sc->pending) { set_error("a frame is pending: call rt_frame_finish first"); return RT_ERR_INVALID_ARG; }
    if ((mode == TR_BVH_BOX || mode == TR_BVH_NOBOX) && !sc->bvh.valid) {
        set_error("rt_cast_rays_static builds nothing: call rt_bvh_build (or render a frame) first");
        return RT_ERR_NO_BVH;
    }
    if (n == 0) return RT_OK;
    std::lock_guard<std::mutex> lock(g_cast_static_mutex);
    unsigned char* base = nullptr;
    if (int rc = cast_static_address(&base)) return rc;
    if (int rc = rth::update_instance_boxes(sc)) return rc;  // one kernel when instances moved; allocates nothing
    const size_t nn = (size_t)n;
    int* di = (int*)base;
    int* dt = di + nn;
    float* dtt = (float*)(di + 2 * nn);
    float* d_o = dtt + nn;
    float* d_d = d_o + 3 * nn;
    cudaStream_t st = sc->stream;
    RT_CUDA(cudaMemcpyAsync(d_o, origins, 12 * nn, cudaMemcpyHostToDevice, st));
    RT_CUDA(cudaMemcpyAsync(d_d, dirs, 12 * nn, cudaMemcpyHostToDevice, st));
    SceneView v = rth::make_view(sc, true);
    const int g = (n + kThreads - 1) / kThreads;
    switch (mode) {
        case TR_BVH_BOX: k_cast_rays<TR_BVH_BOX><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
        case TR_BVH_NOBOX: k_cast_rays<TR_BVH_NOBOX><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
        case TR_BRUTE: k_cast_rays<TR_BRUTE><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
        default: k_cast_rays<TR_BRUTE_BOX><<<g, kThreads, 0, st>>>(v, d_o, d_d, n, di, dt, dtt); break;
    }
    sc->kernel_launches++;
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(instance, di, 4 * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(triangle, dt, 4 * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(t, dtt, 4 * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}
