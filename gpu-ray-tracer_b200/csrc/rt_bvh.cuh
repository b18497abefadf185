// rt_bvh.cuh — compressed wide BVH8: node format and the conservative traversal.
//
// Replaces the reference's implicit binary tree over instance boxes and its stackless,
// warp-uniform iterator (include/rayopt/gpu/bvh.h:26-100, src/rayopt/bvh.cu:98-156) and the
// 12-triangle brute force at every leaf (src/rayprimitives/trimesh.cu:11-19).
//
// The tree is built over WORLD-SPACE triangles (instances x mesh triangles). Traversal is a
// conservative filter: node boxes are inflated supersets, the triangle test accepts a margin
// around every triangle, culling distances carry slack. Whatever survives is handed to
// exact_candidate() (rt_scene.cuh), which decides with the reference's own arithmetic, so the
// answer is the reference's answer and the BVH only removes work.
//
// Node = 80 bytes = 5 x 128-bit loads:
//   q0: origin.x, origin.y, origin.z (fp32), [ex | ey<<8 | ez<<16 | imask<<24]
//       e* = biased fp32 exponent of the per-axis quantisation step, imask = slots holding inner nodes
//   q1: child_base (index of first inner child), prim_base (index of first leaf triangle),
//       meta[0..3], meta[4..7]
//       meta: 0 = empty; inner: 0x20 | (24 + slot); leaf: (unary count 1/3/7) << 5 | offset from prim_base
//   q2: qlo_x[0..7], qlo_y[0..7]    q3: qlo_z[0..7], qhi_x[0..7]    q4: qhi_y[0..7], qhi_z[0..7]
//       child box = origin + q * 2^(e-127); lo rounded down, hi rounded up
// Children sit in slots chosen so that slot ^ octant orders them front to back for a ray of that
// octant (Ylitie, Karras, Laine: "Efficient Incoherent Ray Traversal on GPUs Through Compressed
// Wide BVHs", HPG 2017 — the published algorithm; code written for this repository).
//
// Triangle record = 48 bytes = 3 x 128-bit loads, in leaf order:
//   p0: v0.xyz, instance index     p1: (v1-v0).xyz, global triangle index     p2: (v2-v0).xyz, L
//   L = |e1|_1 + |e2|_1, the length scale of the acceptance margin.
#ifndef RTB200_BVH_CUH
#define RTB200_BVH_CUH

#include <string.h>

#include "rt_scene.cuh"

namespace rtd {

#define RT_STACK_SIZE 32
// absolute slack on culling distances: t values of the exact test differ from geometric
// distances by rounding of the local-frame transforms only (relative ~1e-6)
#define RT_TCULL_REL 1.0001f
#define RT_TCULL_ABS 1e-4f
#ifndef RT_POSTPONE_PCT
#define RT_POSTPONE_PCT 0   /* measured: no gain on coherent primary / shadow rays (profiles/r1_summary.md) */
#endif

RT_HD int hd_clz(unsigned x) {
#ifdef __CUDA_ARCH__
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}
RT_HD int hd_ctz(unsigned x) {
#ifdef __CUDA_ARCH__
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
RT_HD int hd_popc(unsigned x) {
#ifdef __CUDA_ARCH__
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
RT_HD float hd_u2f(unsigned u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
RT_HD int hd_f2i(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_int(f);
#else
    int i; memcpy(&i, &f, 4); return i;
#endif
}
RT_HD uint4 ldg4u(const uint4* p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}
RT_HD float4 ldg4f(const float4* p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}

struct TravStats { unsigned nodes, prims, exact; };

// Conservative ray/triangle filter (division-free Moeller-Trumbore on unnormalised quantities).
// Passes every triangle the reference's unsigned-area test could accept: the acceptance region
// is widened by D = 1e-4 + 2e-5*L + 2e-6*|o - v0|_1 world units around the triangle, which
// covers the reference's 1e-5 barycentric tolerance and the rounding of both computations.
RT_HD bool tri_filter(float3 o, float3 d, float4 p0, float4 p1, float4 p2, float tcull) {
    float3 e1 = rtx::mk3(p1.x, p1.y, p1.z), e2 = rtx::mk3(p2.x, p2.y, p2.z);
    float3 tv = rtx::mk3(o.x - p0.x, o.y - p0.y, o.z - p0.z);
    float3 pv = rtx::mk3(fmaf(d.y, e2.z, -d.z * e2.y), fmaf(d.z, e2.x, -d.x * e2.z), fmaf(d.x, e2.y, -d.y * e2.x));
    float det = fmaf(e1.x, pv.x, fmaf(e1.y, pv.y, e1.z * pv.z));
    float up = fmaf(tv.x, pv.x, fmaf(tv.y, pv.y, tv.z * pv.z));
    float3 qv = rtx::mk3(fmaf(tv.y, e1.z, -tv.z * e1.y), fmaf(tv.z, e1.x, -tv.x * e1.z), fmaf(tv.x, e1.y, -tv.y * e1.x));
    float vp = fmaf(d.x, qv.x, fmaf(d.y, qv.y, d.z * qv.z));
    float tp = fmaf(e2.x, qv.x, fmaf(e2.y, qv.y, e2.z * qv.z));
    float a = fabsf(det);
    if (det < 0.0f) { up = -up; vp = -vp; tp = -tp; }
    float L = p2.w;
    float tv1 = fabsf(tv.x) + fabsf(tv.y) + fabsf(tv.z);
    float tau = (1e-4f + 2e-5f * L + 2e-6f * tv1) * L;
    // t' = e2 . (tvec x e1) carries one more edge factor than u', v'
    float tau_t = fmaf(1e-4f, a, tau * (1.0f + L));
    return up >= -tau && vp >= -tau && up + vp <= a + tau && tp >= -tau_t && tp <= fmaf(fminf(tcull, 3e38f), a, tau_t);
}

// The same filter with a second verdict for the opt-in tolerance mode (SceneView::fast_accept, RT_RENDER_FAST_ACCEPT):
//   0 = reject, 1 = candidate (the reference-exact test decides), 2 = CLEAR hit: the ray crosses the triangle's plane at
//   t >= 1e-4 (ten thresholds), not grazing (|d.n| >= 4e-5 against the reference's 1e-5), and inside all three edges by twice
//   the margin that covers the reference's 1e-5 barycentric tolerance. The reference accepts every such hit; t = t'/|det|,
//   uv = (u', v')/|det| then differ from its values only by rounding (measured: profiles/r2_fast_accept_parity.md).
RT_HD int tri_filter_clear(float3 o, float3 d, float4 p0, float4 p1, float4 p2, float tcull, float& t, float& u, float& v) {
    float3 e1 = rtx::mk3(p1.x, p1.y, p1.z), e2 = rtx::mk3(p2.x, p2.y, p2.z);
    float3 tv = rtx::mk3(o.x - p0.x, o.y - p0.y, o.z - p0.z);
    float3 pv = rtx::mk3(fmaf(d.y, e2.z, -d.z * e2.y), fmaf(d.z, e2.x, -d.x * e2.z), fmaf(d.x, e2.y, -d.y * e2.x));
    float det = fmaf(e1.x, pv.x, fmaf(e1.y, pv.y, e1.z * pv.z));
    float up = fmaf(tv.x, pv.x, fmaf(tv.y, pv.y, tv.z * pv.z));
    float3 qv = rtx::mk3(fmaf(tv.y, e1.z, -tv.z * e1.y), fmaf(tv.z, e1.x, -tv.x * e1.z), fmaf(tv.x, e1.y, -tv.y * e1.x));
    float vp = fmaf(d.x, qv.x, fmaf(d.y, qv.y, d.z * qv.z));
    float tp = fmaf(e2.x, qv.x, fmaf(e2.y, qv.y, e2.z * qv.z));
    float a = fabsf(det);
    if (det < 0.0f) { up = -up; vp = -vp; tp = -tp; }
    float L = p2.w;
    float tv1 = fabsf(tv.x) + fabsf(tv.y) + fabsf(tv.z);
    float tau = (1e-4f + 2e-5f * L + 2e-6f * tv1) * L;
    float tau_t = fmaf(1e-4f, a, tau * (1.0f + L));
    if (!(up >= -tau && vp >= -tau && up + vp <= a + tau && tp >= -tau_t && tp <= fmaf(fminf(tcull, 3e38f), a, tau_t))) return 0;
    // clear? (L = +inf marks an analytic sphere: never clear)
    if (up >= 2.0f * tau && vp >= 2.0f * tau && up + vp <= a - 2.0f * tau && a >= 1e-5f * L * L && tp >= fmaf(1e-4f, a, tau_t) && L < 3e38f) {
        float inv = 1.0f / a;
        t = tp * inv; u = up * inv; v = vp * inv;
        return 2;
    }
    return 1;
}

struct RayPrep {
    float3 o, d, idir;
    unsigned octinv;   // 7 - octant
    unsigned octinv4;  // octinv replicated into 4 bytes
};

RT_HD RayPrep ray_prep(const rtx::Ray& r) {
    RayPrep p;
    p.o = r.o; p.d = r.d;
    const float eps = 5.4210108624275222e-20f;  // 2^-64: keeps (lo - o) * idir finite
    float dx = fabsf(r.d.x) > eps ? r.d.x : copysignf(eps, r.d.x);
    float dy = fabsf(r.d.y) > eps ? r.d.y : copysignf(eps, r.d.y);
    float dz = fabsf(r.d.z) > eps ? r.d.z : copysignf(eps, r.d.z);
    p.idir = rtx::mk3(1.0f / dx, 1.0f / dy, 1.0f / dz);
    unsigned oct = (dx < 0.0f ? 1u : 0u) | (dy < 0.0f ? 2u : 0u) | (dz < 0.0f ? 4u : 0u);
    p.octinv = 7u - oct;
    p.octinv4 = p.octinv * 0x01010101u;
    return p;
}

RT_HD unsigned byte_of(unsigned w, int i) { return (w >> (8 * i)) & 0xffu; }

// Byte i of w as the float 1 + q * 2^-15 (q in mantissa bits 8..15): one PRMT on the device, no
// integer-to-float conversion (I2F issues at a quarter of the FFMA rate and was the co-bottleneck
// of the node test). With A = 2^15 * a and B = b - A, fma(f, A, B) = q * a + b.
// `one` must hold 0x3f800000 where the compiler cannot fold it (SceneView::one_bits, a kernel
// parameter): PRMT takes
// a single immediate, and with the constant as the immediate the selector was re-materialised from
// a uniform register for every byte (55 extra MOV/IMAD per node in the first version).
template <int I>
RT_HD float qfloat(unsigned w, unsigned one) {
#ifdef __CUDA_ARCH__
    unsigned r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(one), "n"(0x7604 | (I << 4)));
    return __uint_as_float(r);
#else
    return hd_u2f(one | (byte_of(w, I) << 8));
#endif
}
// byte I of w, zero-extended: one PRMT with an immediate selector
template <int I>
RT_HD unsigned qbyte(unsigned w) {
#ifdef __CUDA_ARCH__
    unsigned r;
    asm("prmt.b32 %0, %1, 0, %2;" : "=r"(r) : "r"(w), "n"(0x4440 | I));
    return r;
#else
    return byte_of(w, I);
#endif
}

// Tests the 8 quantised child boxes of one node. Returns the hit mask in traversal layout:
// bits 24..31 inner children at position 24 + (slot ^ octinv), bits 0..23 leaf triangles.
// Conservative: the entry distance is biased down and the exit distance up by more than the
// rounding of the fused multiply-adds, so a box the exact slab test would hit is never missed.
RT_HD unsigned node_children_hits(const uint4 q0, const uint4 q1, const uint4 q2, const uint4 q3, const uint4 q4,
                                  const RayPrep& rp, float tcull, const unsigned one) {
    float px = hd_u2f(q0.x), py = hd_u2f(q0.y), pz = hd_u2f(q0.z);
    // 2^15 * 2^(e-127): exponent byte + 15
    float sx = hd_u2f(((q0.w & 0xffu) + 15u) << 23), sy = hd_u2f((((q0.w >> 8) & 0xffu) + 15u) << 23),
          sz = hd_u2f((((q0.w >> 16) & 0xffu) + 15u) << 23);
    float Ax = sx * rp.idir.x, Ay = sy * rp.idir.y, Az = sz * rp.idir.z;
    float bx = (px - rp.o.x) * rp.idir.x, by = (py - rp.o.y) * rp.idir.y, bz = (pz - rp.o.z) * rp.idir.z;
    float Bx = bx - Ax, By = by - Ay, Bz = bz - Az;
    const float kSlack = 2.384185791015625e-07f;  // 2^-22
    float dx = (fabsf(Ax) + fabsf(bx)) * kSlack, dy = (fabsf(Ay) + fabsf(by)) * kSlack, dz = (fabsf(Az) + fabsf(bz)) * kSlack;
    float B0x = Bx - dx, B1x = Bx + dx, B0y = By - dy, B1y = By + dy, B0z = Bz - dz, B1z = Bz + dz;
    bool nx = rp.idir.x < 0.0f, ny = rp.idir.y < 0.0f, nz = rp.idir.z < 0.0f;
    unsigned hits = 0;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        unsigned meta4 = half ? q1.w : q1.z;
        unsigned lox = half ? q2.y : q2.x, loy = half ? q2.w : q2.z, loz = half ? q3.y : q3.x;
        unsigned hix = half ? q3.w : q3.z, hiy = half ? q4.y : q4.x, hiz = half ? q4.w : q4.z;
        unsigned nearx = nx ? hix : lox, farx = nx ? lox : hix;
        unsigned neary = ny ? hiy : loy, fary = ny ? loy : hiy;
        unsigned nearz = nz ? hiz : loz, farz = nz ? loz : hiz;
        // four children at once: inner slots (offsets 24..31) get their position XORed with the
        // ray octant, leaves keep their triangle offset; empty slots have zero child bits
        unsigned is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        unsigned inner_mask4 = (is_inner4 >> 4) * 0xffu;
        unsigned bit_index4 = (meta4 ^ (rp.octinv4 & inner_mask4)) & 0x1f1f1f1fu;
        unsigned child_bits4 = (meta4 >> 5) & 0x07070707u;
#define RT_CHILD(I)                                                                                          \
        {                                                                                                    \
            float t0x = fmaf(qfloat<I>(nearx, one), Ax, B0x), t1x = fmaf(qfloat<I>(farx, one), Ax, B1x);      \
            float t0y = fmaf(qfloat<I>(neary, one), Ay, B0y), t1y = fmaf(qfloat<I>(fary, one), Ay, B1y);      \
            float t0z = fmaf(qfloat<I>(nearz, one), Az, B0z), t1z = fmaf(qfloat<I>(farz, one), Az, B1z);      \
            float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));                                           \
            float tmax = fminf(fminf(t1x, t1y), fminf(t1z, tcull));                                          \
            unsigned bits = qbyte<I>(child_bits4) << qbyte<I>(bit_index4);                                   \
            if (tmin <= tmax) hits |= bits;                                                                  \
        }
        RT_CHILD(0) RT_CHILD(1) RT_CHILD(2) RT_CHILD(3)
#undef RT_CHILD
    }
    return hits;
}

// Closest-hit (ANY == false) or occlusion (ANY == true: stop at the first accepted hit with
// t <= max_t) query. `use_box` applies the reference's leaf-box condition to candidates.
// FAST: the opt-in tolerance mode (tri_filter_clear) as a compile-time variant, so that the default kernels carry none of it
// (as a run-time branch it cost the bit-exact path 2.5 %).
template <bool ANY, bool FAST = false>
RT_HD bool bvh_query(const SceneView& s, const rtx::Ray& r, bool use_box, float max_t, Hit& best, TravStats* st) {
    hit_reset(best);
    if (s.n_prims <= 0) return false;
    RayPrep rp = ray_prep(r);
    InstCtx ctx;
    ctx.inst = -1;
    float tcull = ANY ? fmaf(max_t, RT_TCULL_REL, RT_TCULL_ABS) : INFINITY;
    uint2 stack[RT_STACK_SIZE];
    int sp = 0;
    uint2 ng;  // node group: x = base index, y = hit bits 24..31 | imask
    ng.x = 0u; ng.y = 0x80000000u;
    for (;;) {
        uint2 tg;  // triangle group: x = prim base, y = 24 hit bits
        if (ng.y > 0x00ffffffu) {
            unsigned hits = ng.y;
            unsigned bit = 31u - (unsigned)hd_clz(hits);
            hits &= ~(1u << bit);
            if (hits & 0xff000000u) {
                uint2 e; e.x = ng.x; e.y = hits;
                if (sp < RT_STACK_SIZE) stack[sp++] = e;
            }
            unsigned slot = (bit - 24u) ^ rp.octinv;
            unsigned rel = (unsigned)hd_popc(hits & 0xffu & ((1u << slot) - 1u));
            const uint4* np = s.nodes + (size_t)(ng.x + rel) * 5;
            uint4 q0 = ldg4u(np), q1 = ldg4u(np + 1), q2 = ldg4u(np + 2), q3 = ldg4u(np + 3), q4 = ldg4u(np + 4);
            if (st) st->nodes++;
            unsigned h = node_children_hits(q0, q1, q2, q3, q4, rp, tcull, s.one_bits);
            ng.x = q1.x; ng.y = (h & 0xff000000u) | (q0.w >> 24);
            tg.x = q1.y; tg.y = h & 0x00ffffffu;
        } else {
            tg = ng;
            ng.x = 0u; ng.y = 0u;
        }
#ifdef __CUDA_ARCH__
        // Triangle postponing (Ylitie et al. 2017, §4.3): when only a few lanes of the warp have
        // triangles to test and this lane still has nodes to visit, park the triangle group on the
        // stack and keep traversing, so that triangle tests run with more lanes active. The order
        // of tests does not change the result (closest hit with index tie rules / any hit).
        if (RT_POSTPONE_PCT > 0 && tg.y != 0u && ng.y > 0x00ffffffu && sp < RT_STACK_SIZE - 2) {
            unsigned act = __activemask();
            int with_tris = __popc(__ballot_sync(act, tg.y != 0u));
            if (with_tris * 100 < __popc(act) * RT_POSTPONE_PCT) {
                stack[sp++] = tg;
                tg.y = 0u;
            }
        }
#endif
        while (tg.y) {
            unsigned k = (unsigned)hd_ctz(tg.y);
            tg.y &= tg.y - 1u;
            const float4* pp = s.prims + (size_t)(tg.x + k) * 3;
            float4 p0 = ldg4f(pp), p1 = ldg4f(pp + 1), p2 = ldg4f(pp + 2);
            if (st) st->prims++;
            float ft, fu, fv;
            const int verdict = FAST ? tri_filter_clear(rp.o, rp.d, p0, p1, p2, tcull, ft, fu, fv)
                                     : (tri_filter(rp.o, rp.d, p0, p1, p2, tcull) ? 1 : 0);
            if (verdict == 2 && (!ANY || ft <= max_t * 0.9999f)) {
                // tolerance mode: a clear hit is taken as it is (fast_candidate applies the same ordering rules)
                if (fast_candidate(hd_f2i(p0.w), hd_f2i(p1.w), ft, fu, fv, best)) {
                    if (ANY) return true;
                    tcull = fmaf(best.t, RT_TCULL_REL, RT_TCULL_ABS);
                }
            } else if (verdict) {
                if (st) st->exact++;
                int inst = hd_f2i(p0.w), tri = hd_f2i(p1.w);
                if (exact_candidate(s, r, inst, tri, use_box, ctx, best)) {
                    if (ANY) {
                        if (!(best.t > max_t)) return true;
                        hit_reset(best);  // accepted but beyond the light: keep looking
                    } else {
                        tcull = fmaf(best.t, RT_TCULL_REL, RT_TCULL_ABS);
                    }
                }
            }
        }
        if (ng.y <= 0x00ffffffu) {
            if (sp == 0) break;
            ng = stack[--sp];
        }
    }
    return ANY ? false : best.inst >= 0;
}

template <bool FAST>
struct BvhTracerT {
    const SceneView* s;
    bool use_box;
    TravStats* st;
    RT_HD bool closest(const rtx::Ray& r, Hit& best) const { return bvh_query<false, FAST>(*s, r, use_box, INFINITY, best, st); }
    RT_HD bool occluded(const rtx::Ray& r, float max_t) const {
        Hit h;
        return bvh_query<true, FAST>(*s, r, use_box, max_t, h, st);
    }
};
typedef BvhTracerT<false> BvhTracer;

}  // namespace rtd
#endif
