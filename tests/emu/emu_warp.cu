// emu_warp.cu — TEST INFRASTRUCTURE: the warp-coherent traversal of the CUDA engine (csrc/rt_beam.cuh, __device__ code
// written with warp intrinsics) executed on the CPU, unmodified.
//
// rt_beam.cuh walks the 32 rays of a warp together: shuffles, ballots, a warp-wide OR, one stack spread over the lanes'
// registers. There is no GPU in the build container, so this harness gives that header a warp: 32 fibers (ucontext) per
// emulated warp, scheduled round-robin on one OS thread; a warp collective is a rendezvous — every fiber deposits its
// value and yields until the last one has arrived — which is exactly the *_sync contract (all 32 lanes call it, full mask).
// The header is included as it is; for the host pass only, the device intrinsics it uses are mapped to the stand-ins below
// and `__device__` is widened to `__host__ __device__`. Nothing here is a product path (librtb200.so contains none of it).
// Built by __graft_entry__.build() into tests/emu/libemu.so next to emu.cu.
#include <ucontext.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

#include "rt_bvh.cuh"
#include "rt_hostprep.h"
#include "rt_shade.cuh"

#ifndef __CUDA_ARCH__
namespace emuw {

struct Warp {
    ucontext_t sched;
    ucontext_t ctx[32];
    std::vector<char> stack[32];
    bool done[32];
    int cur = 0;
    unsigned buf[2][32];
    int arrived = 0;
    unsigned gen = 0;
    std::function<void(int)> body;
};
static thread_local Warp* g_warp = nullptr;

struct Tid { unsigned x, y, z; };
static inline Tid tid() { Tid t = {(unsigned)g_warp->cur, 0u, 0u}; return t; }

static inline void yield_lane() {
    Warp* w = g_warp;
    swapcontext(&w->ctx[w->cur], &w->sched);
}
// every lane deposits v; returns the 32 values (valid until the collective after the next one: two buffers alternate)
static inline const unsigned* rendezvous(unsigned v) {
    Warp* w = g_warp;
    const unsigned g = w->gen;
    unsigned* b = w->buf[g & 1u];
    b[w->cur] = v;
    if (++w->arrived == 32) { w->arrived = 0; w->gen = g + 1u; }
    else {
        long spins = 0;
        while (w->gen == g) {
            yield_lane();
            if (++spins > 100000) { fprintf(stderr, "emu_warp: lane %d waits for a collective the other lanes never reach\n", w->cur); abort(); }
        }
    }
    return b;
}
template <class T> static inline unsigned bits_of(T v) { static_assert(sizeof(T) == 4, "32-bit values only"); unsigned u; memcpy(&u, &v, 4); return u; }
template <class T> static inline T from_bits(unsigned u) { T v; memcpy(&v, &u, 4); return v; }

template <class T> static inline T shfl(unsigned, T v, int src) { const unsigned* b = rendezvous(bits_of(v)); return from_bits<T>(b[src & 31]); }
template <class T> static inline T shfl_xor(unsigned, T v, int m) { const int l = g_warp->cur; const unsigned* b = rendezvous(bits_of(v)); return from_bits<T>(b[(l ^ m) & 31]); }
static inline unsigned ballot(unsigned, int p) {
    const unsigned* b = rendezvous(p ? 1u : 0u);
    unsigned m = 0;
    for (int i = 0; i < 32; i++) m |= b[i] << i;
    return m;
}
static inline int any(unsigned m, int p) { return ballot(m, p) != 0u; }
static inline unsigned reduce_or(unsigned, unsigned v) {
    const unsigned* b = rendezvous(v);
    unsigned m = 0;
    for (int i = 0; i < 32; i++) m |= b[i];
    return m;
}
static inline int ffs_(int x) { return x ? __builtin_ctz((unsigned)x) + 1 : 0; }
static inline int clz_(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline int popc_(unsigned x) { return __builtin_popcount(x); }
// PTX prmt, default mode: selector nibble = source byte 0..7 of {y:x}, bit 3 = replicate that byte's sign bit
static inline unsigned byte_perm(unsigned x, unsigned y, unsigned s) {
    const unsigned long long src = ((unsigned long long)y << 32) | x;
    unsigned r = 0;
    for (int i = 0; i < 4; i++) {
        const unsigned sel = (s >> (4 * i)) & 0xfu;
        unsigned byte = (unsigned)(src >> (8 * (sel & 7u))) & 0xffu;
        if (sel & 8u) byte = (byte & 0x80u) ? 0xffu : 0x00u;
        r |= byte << (8 * i);
    }
    return r;
}
static inline float u2f(unsigned u) { return from_bits<float>(u); }
static inline int f2i(float f) { return (int)bits_of(f); }

static void fiber_entry() {
    Warp* w = g_warp;
    const int lane = w->cur;
    w->body(lane);
    w->done[lane] = true;
    for (;;) yield_lane();  // never resumed once done
}

// Runs body(lane) on 32 fibers until all have returned.
static void run_warp(Warp& w, std::function<void(int)> body) {
    w.body = std::move(body);
    w.arrived = 0; w.gen = 0;
    g_warp = &w;
    for (int i = 0; i < 32; i++) {
        w.done[i] = false;
        if (w.stack[i].empty()) w.stack[i].resize(512 * 1024);
        getcontext(&w.ctx[i]);
        w.ctx[i].uc_stack.ss_sp = w.stack[i].data();
        w.ctx[i].uc_stack.ss_size = w.stack[i].size();
        w.ctx[i].uc_link = nullptr;
        makecontext(&w.ctx[i], fiber_entry, 0);
    }
    for (;;) {
        bool any_alive = false;
        for (int i = 0; i < 32; i++) {
            if (w.done[i]) continue;
            any_alive = true;
            w.cur = i;
            swapcontext(&w.sched, &w.ctx[i]);
        }
        if (!any_alive) break;
    }
    g_warp = nullptr;
}

}  // namespace emuw

// ---- for the host pass of rt_beam.cuh only
#define threadIdx (emuw::tid())
#define __shfl_sync emuw::shfl
#define __shfl_xor_sync emuw::shfl_xor
#define __ballot_sync emuw::ballot
#define __any_sync emuw::any
#define __reduce_or_sync emuw::reduce_or
#define __ffs emuw::ffs_
#define __clz emuw::clz_
#define __popc emuw::popc_
#define __byte_perm emuw::byte_perm
#define __uint_as_float emuw::u2f
#define __float_as_int emuw::f2i
#undef __device__
#define __device__ __location__(host) __location__(device)
#endif  // !__CUDA_ARCH__

#include "rt_beam.cuh"

#ifndef __CUDA_ARCH__
#undef __device__
#define __device__ __location__(device)
#undef threadIdx

using namespace rtd;

namespace {

struct WScene {
    rth::HostScene hs;
    std::vector<InstRec> insts;
    SceneView v;
};

void fill_view(WScene* e, const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims) {
    rth::prepare_host_scene(d, e->hs);
    e->insts = rth::prepare_instances(d->instances, d->n_instances);
    for (auto& I : e->insts) instance_box(I, e->hs.meshes.data(), e->hs.mesh_box.data());
    SceneView& v = e->v;
    std::memset(&v, 0, sizeof v);
    v.verts = e->hs.verts.data(); v.norms = e->hs.norms.data(); v.tris = e->hs.tris.data();
    v.meshes = e->hs.meshes.data(); v.insts = e->insts.data(); v.mats = e->hs.mats.data(); v.lights = e->hs.lights.data();
    if (!e->hs.texc.empty()) { v.texc = e->hs.texc.data(); v.atlas = e->hs.atlas.data(); v.atlas_w = e->hs.atlas_w; v.atlas_h = e->hs.atlas_h; }
    v.n_insts = d->n_instances; v.n_lights = d->n_lights;
    v.n_prims = n_prims; v.nodes = (const uint4*)nodes8; v.prims = (const float4*)prims8;
    const RtCamera& c = d->camera;
    v.cam = rtx::camera_basis(c.pose.rot, c.pose.pos, c.global_near, c.unit_to_pixels, c.canvas_width, c.canvas_height);
    v.ambience = make_float4(d->env.ambience[0], d->env.ambience[1], d->env.ambience[2], d->env.ambience[3]);
    v.dist_c0 = d->env.dist_atten[0]; v.dist_c1 = d->env.dist_atten[1]; v.dist_c2 = d->env.dist_atten[2];
    v.depth = d->env.depth; v.width = d->width; v.height = d->height;
    v.any_refractive = e->hs.any_refractive; v.any_secondary = e->hs.any_secondary;
    v.one_bits = 0x3f800000u;
}

// tiles of 8x4 pixels, lane l -> pixel (8 tx + (l & 7), 4 ty + (l >> 3)): the work unit of k_tile (rt_engine.cu TileMap::locate)
template <class F>
void for_tiles(int W, int H, int threads, F f) {
    const int tx = (W + 7) / 8, ty = (H + 3) / 4, n = tx * ty;
    threads = std::max(1, std::min(threads, n));
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([=]() {
            emuw::Warp* w = new emuw::Warp();
            emuw::run_warp(*w, [&](int lane) {
                for (int q = t; q < n; q += threads) {
                    const int x = (q % tx) * 8 + (lane & 7), y = (q / tx) * 4 + (lane >> 3);
                    f(lane, x, y, x < W && y < H);
                }
            });
            delete w;
        });
    for (auto& x : th) x.join();
}

}  // namespace

extern "C" {

// Camera rays through beam_query as k_tile<1> calls it (closest hit, leaf-box rule on, tolerance mode off).
// stats4: nodes visited (per active ray), triangle filters, exact tests, node fetches (per warp).
// use_box: 0 = BVH8 only, 1 = with the reference's leaf-box rule, 2 = leaf-box rule + the opt-in tolerance mode
// (beam_query<.., FAST = true>: clear filter hits skip the reference-exact test; scenes without rotations only).
int emu_beam_primary(const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims, int use_box,
                     int32_t* inst, int32_t* tri, float* t, uint64_t* stats4, int threads) {
    WScene e;
    fill_view(&e, d, nodes8, prims8, n_prims);
    const SceneView& v = e.v;
    uint64_t tot[4] = {0, 0, 0, 0};
    // per-lane counters live on the fibers' stacks; every pixel keeps its own and they are added up at the end
    std::vector<TravStats> row_ts((size_t)v.height * v.width, TravStats{0, 0, 0});
    std::vector<unsigned> row_nf((size_t)v.height * v.width, 0u);
    for_tiles(v.width, v.height, threads, [&](int lane, int x, int y, bool valid) {
        (void)lane;
        rtx::Ray r;
        r.o = v.cam.pos; r.d = v.cam.f;
        if (valid) r = rtx::camera_ray(v.cam, (float)x, (float)y);
        Hit h;
        TravStats ts = {0, 0, 0};
        BeamStats bs = {0};
        const bool hit = use_box == 2 ? beam_query<true, true>(v, r, valid, false, true, INFINITY, h, ts, bs)
                                      : beam_query<true, false>(v, r, valid, false, use_box != 0, INFINITY, h, ts, bs);
        if (valid) {
            const size_t p = (size_t)y * v.width + x;
            inst[p] = hit ? h.inst : -1;
            tri[p] = hit ? h.tri - v.meshes[v.insts[h.inst].mesh].tri_begin : -1;
            t[p] = h.t;
            row_ts[p] = ts;
            row_nf[p] = bs.node_fetches;
        }
    });
    for (size_t p = 0; p < row_ts.size(); p++) { tot[0] += row_ts[p].nodes; tot[1] += row_ts[p].prims; tot[2] += row_ts[p].exact; tot[3] += row_nf[p]; }
    if (stats4) for (int k = 0; k < 4; k++) stats4[k] = tot[k];
    return 0;
}

// Shadow rays of the camera rays' hits towards every light: the occlusion query of k_shadow_beam / k_tile (one walk per
// tile and light) next to the per-lane query of k_shadow (BvhTracer::occluded), both over the same rays.
// occl_beam / occl_lane: n_lights x W x H bytes, 255 where no shadow ray exists (miss, or a light with zero Phong weight
// when cull != 0 — the rays the kernels do not trace either).
int emu_beam_shadow(const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims, int cull,
                    unsigned char* occl_beam, unsigned char* occl_lane, int threads) {
    WScene e;
    fill_view(&e, d, nodes8, prims8, n_prims);
    SceneView& v = e.v;
    v.cull_unlit = (cull && e.hs.cull_ok && !e.hs.any_refractive) ? 1 : 0;
    const size_t plane = (size_t)v.width * v.height;
    std::memset(occl_beam, 255, plane * (size_t)std::max(v.n_lights, 1));
    std::memset(occl_lane, 255, plane * (size_t)std::max(v.n_lights, 1));
    for_tiles(v.width, v.height, threads, [&](int lane, int x, int y, bool valid) {
        (void)lane;
        rtx::Ray r;
        r.o = v.cam.pos; r.d = v.cam.f;
        if (valid) r = rtx::camera_ray(v.cam, (float)x, (float)y);
        Hit h;
        TravStats ts = {0, 0, 0};
        BeamStats bs = {0};
        const bool hit = beam_query<false, false>(v, r, valid, false, true, INFINITY, h, ts, bs);
        float3 nrm = rtx::mk3(0.f, 0.f, 0.f), hit_pos = rtx::mk3(0.f, 0.f, 0.f);
        int mat_id = 0;
        if (hit) {
            mat_id = v.tris[h.tri].mat;
            nrm = hit_normal(v, h, r);
            hit_pos = rtx::ray_at(r, h.t);
        }
        const MatRec& m = v.mats[mat_id];
        for (int l = 0; l < v.n_lights; l++) {
            bool trace = false;
            rtx::Ray sr;
            sr.o = hit_pos; sr.d = r.d;
            float3 dir = r.d;
            float max_t = 0.0f, da = 1.0f;
            if (hit) {
                light_setup(v, v.lights[l], hit_pos, dir, max_t, da);
                if (!(v.cull_unlit && light_unlit(m, nrm, dir))) {
                    trace = true;
                    sr = shadow_ray(hit_pos, dir);
                }
            }
            Hit sh;
            const bool blocked = beam_query<false, false>(v, sr, trace, true, true, max_t, sh, ts, bs);
            if (trace) {
                const size_t j = (size_t)l * plane + (size_t)y * v.width + x;
                occl_beam[j] = blocked ? 1 : 0;
                BvhTracer tr; tr.s = &v; tr.use_box = true; tr.st = nullptr;
                occl_lane[j] = tr.occluded(sr, max_t) ? 1 : 0;
            }
        }
    });
    return 0;
}

// Arbitrary rays, 32 per warp in the order given (closest hit): rays after a bounce, mixed octants -> the per-lane fallback.
int emu_beam_cast_rays(const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims, const float* origins,
                       const float* dirs, int n, int32_t* inst, int32_t* tri, float* t) {
    WScene e;
    fill_view(&e, d, nodes8, prims8, n_prims);
    const SceneView& v = e.v;
    emuw::Warp* w = new emuw::Warp();
    emuw::run_warp(*w, [&](int lane) {
        for (int base = 0; base < n; base += 32) {
            const int i = base + lane;
            const bool valid = i < n;
            rtx::Ray r;
            r.o = v.cam.pos; r.d = v.cam.f;
            if (valid) r = rtx::ray_make(rtx::mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                                         rtx::mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
            Hit h;
            TravStats ts = {0, 0, 0};
            BeamStats bs = {0};
            const bool hit = beam_query<false, false>(v, r, valid, false, true, INFINITY, h, ts, bs);
            if (valid) {
                inst[i] = hit ? h.inst : -1;
                tri[i] = hit ? h.tri - v.meshes[v.insts[h.inst].mesh].tri_begin : -1;
                t[i] = h.t;
            }
        }
    });
    delete w;
    return 0;
}

// Self-test of the stand-ins: prmt selector semantics and a shuffle / ballot / reduce round on one warp. Returns mismatches.
int emu_warp_selftest(void) {
    int bad = 0;
    if (emuw::byte_perm(0x03020100u, 0x07060504u, 0x00000005u) != 0x00000005u) bad++;
    if (emuw::byte_perm(0x03020100u, 0x07060504u, 0x00007531u) != 0x07050301u) bad++;
    if (emuw::byte_perm(0x80000000u, 0u, 0x0000000bu) != 0x000000ffu) bad++;
    if (emuw::ffs_(0) != 0 || emuw::ffs_(8) != 4 || emuw::clz_(0) != 32 || emuw::clz_(1) != 31 || emuw::popc_(0xf0f0u) != 8) bad++;
    int lane_bad[32] = {0};
    emuw::Warp* w = new emuw::Warp();
    emuw::run_warp(*w, [&](int lane) {
        for (int round = 0; round < 50; round++) {
            const unsigned v = (unsigned)(lane * 7 + round);
            if (emuw::shfl(0xffffffffu, v, (lane + round) & 31) != (unsigned)(((lane + round) & 31) * 7 + round)) lane_bad[lane]++;
            if (emuw::shfl_xor(0xffffffffu, (float)lane, 5) != (float)(lane ^ 5)) lane_bad[lane]++;
            if (emuw::ballot(0xffffffffu, (lane + round) & 1) != (((round & 1) ? 0x55555555u : 0xaaaaaaaau))) lane_bad[lane]++;
            if (emuw::reduce_or(0xffffffffu, 1u << lane) != 0xffffffffu) lane_bad[lane]++;
            if (emuw::any(0xffffffffu, lane == round % 32) != 1) lane_bad[lane]++;
            if ((lane & 3) == (round & 3)) { volatile int spin = 0; for (int k = 0; k < 10; k++) spin += k; }  // lanes drift apart between collectives
        }
    });
    delete w;
    for (int i = 0; i < 32; i++) bad += lane_bad[i];
    return bad;
}

}  // extern "C"
#endif  // !__CUDA_ARCH__
