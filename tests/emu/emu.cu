// emu.cu — TEST INFRASTRUCTURE: host-side execution of the engine's __host__ __device__ code.
//
// There is no GPU in the build container, so the traversal / exact-test / shading headers of the
// CUDA engine are written __host__ __device__ and this harness runs them on the CPU over a BVH8
// produced by the oracle's host build (oracle/lbvh_host.c). It exists to catch logic errors
// before a GPU run; it is NOT a product path (librtb200.so contains none of this and fails
// loudly without a GPU). Built by __graft_entry__.build() into tests/emu/libemu.so.
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "rt_bvh.cuh"
#include "rt_hostprep.h"
#include "rt_shade.cuh"

using namespace rtd;

struct EmuScene {
    rth::HostScene hs;
    std::vector<InstRec> insts;
    SceneView v;
};

static void fill_view(EmuScene* e, const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims) {
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

template <class F>
static void parallel_rows(int h, int threads, F f) {
    if (threads < 1) threads = 1;
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([=]() { for (int y = t; y < h; y += threads) f(y); });
    for (auto& x : th) x.join();
}

extern "C" {

// mode: 0 brute, 1 BVH8 + leaf-box rule, 2 BVH8 only, 3 brute + leaf-box rule
int emu_trace_primary(const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims, int mode,
                      int32_t* inst, int32_t* tri, float* t, uint64_t* stats3, int threads) {
    EmuScene e;
    fill_view(&e, d, nodes8, prims8, n_prims);
    const SceneView& v = e.v;
    std::vector<TravStats> st((size_t)v.height);
    parallel_rows(v.height, threads, [&](int y) {
        TravStats ts = {0, 0, 0};
        for (int x = 0; x < v.width; x++) {
            rtx::Ray r = rtx::camera_ray(v.cam, (float)x, (float)y);
            Hit h;
            bool hit;
            if (mode == 1 || mode == 2) { BvhTracer tr; tr.s = &v; tr.use_box = mode == 1; tr.st = &ts; hit = tr.closest(r, h); }
            else { BruteTracer tr; tr.s = &v; tr.use_box = mode == 3; hit = tr.closest(r, h); }
            size_t p = (size_t)y * v.width + x;
            inst[p] = hit ? h.inst : -1;
            tri[p] = hit ? h.tri - v.meshes[v.insts[h.inst].mesh].tri_begin : -1;
            t[p] = h.t;
        }
        st[(size_t)y] = ts;
    });
    if (stats3) {
        stats3[0] = stats3[1] = stats3[2] = 0;
        for (auto& s : st) { stats3[0] += s.nodes; stats3[1] += s.prims; stats3[2] += s.exact; }
    }
    return 0;
}

int emu_cast_rays(const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims, int mode,
                  const float* origins, const float* dirs, int n, int32_t* inst, int32_t* tri, float* t) {
    EmuScene e;
    fill_view(&e, d, nodes8, prims8, n_prims);
    const SceneView& v = e.v;
    for (int i = 0; i < n; i++) {
        rtx::Ray r = rtx::ray_make(rtx::mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                                   rtx::mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
        Hit h;
        bool hit;
        if (mode == 1 || mode == 2) { BvhTracer tr; tr.s = &v; tr.use_box = mode == 1; tr.st = nullptr; hit = tr.closest(r, h); }
        else { BruteTracer tr; tr.s = &v; tr.use_box = mode == 3; hit = tr.closest(r, h); }
        inst[i] = hit ? h.inst : -1;
        tri[i] = hit ? h.tri - v.meshes[v.insts[h.inst].mesh].tri_begin : -1;
        t[i] = h.t;
    }
    return 0;
}

// Full frame through propagate_gpu / propagate_cpu_rec + pack_color. rays3: primary, shadow, secondary.
int emu_render(const RtSceneDesc* d, const uint32_t* nodes8, const float* prims8, int n_prims, int mode, int semantics,
               uint32_t* canvas, uint64_t* rays3, int threads) {
    EmuScene e;
    fill_view(&e, d, nodes8, prims8, n_prims);
    const SceneView& v = e.v;
    std::vector<RayCounts> rcs((size_t)v.height);
    parallel_rows(v.height, threads, [&](int y) {
        RayCounts rc = {0, 0, 0};
        for (int x = 0; x < v.width; x++) {
            rtx::Ray r = rtx::camera_ray(v.cam, (float)x, (float)y);
            float4 c;
            auto run = [&](auto tr) {
                if (semantics == RT_SEMANTICS_CPU) {
                    CpuIsect is; is.time = INFINITY; is.norm = rtx::mk3(0.f, 0.f, 0.f); is.mat = 0;
                    float last = 0.f;
                    return propagate_cpu_rec(v, tr, r, is, last, 0, v.depth, 0, rc);
                }
                return propagate_gpu(v, tr, r, rc);
            };
            if (mode == 1 || mode == 2) { BvhTracer tr; tr.s = &v; tr.use_box = mode == 1; tr.st = nullptr; c = run(tr); }
            else { BruteTracer tr; tr.s = &v; tr.use_box = mode == 3; c = run(tr); }
            canvas[(size_t)y * v.width + x] = rtx::pack_color(c);
        }
        rcs[(size_t)y] = rc;
    });
    if (rays3) {
        rays3[0] = rays3[1] = rays3[2] = 0;
        for (auto& r : rcs) { rays3[0] += r.primary; rays3[1] += r.shadow; rays3[2] += r.secondary; }
    }
    return 0;
}

// q == identity fast path against the general quaternion sandwich, bit for bit. Returns mismatches.
int emu_check_identity_rotation(const float* vecs, int n) {
    rtx::Quat q = {0.f, 0.f, 0.f, 1.f};
    int bad = 0;
    for (int i = 0; i < n; i++) {
        float3 v = rtx::mk3(vecs[3 * i], vecs[3 * i + 1], vecs[3 * i + 2]);
        float3 a = rtx::qrotate(q, v), b = rtx::qrotate_identity(v);
        float3 c = rtx::qrotate(rtx::qinverse(q), v);
        if (!(a.x == b.x && a.y == b.y && a.z == b.z && c.x == b.x && c.y == b.y && c.z == b.z)) bad++;
    }
    return bad;
}

}  // extern "C"
