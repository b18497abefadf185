// rt_beam.cuh — warp-coherent BVH8 traversal: one walk per 8x4-pixel tile instead of 32.
//
// Replaces the warp-union walk of the reference (every lane of a warp follows the union of the warp's paths through the
// instance tree, src/rayenv/scene.cu:54-69 + src/rayopt/bvh.cu:98-156) and, for the rays of a screen tile, the per-lane
// walk of rt_bvh.cuh (bvh_query): there every lane decodes and tests the 8 children of every node for its own ray —
// 279 SASS instructions per node visit, executed 32-wide for nodes the whole tile shares (4.78 warp-level visits against
// 4.69 per ray on world16 at 4K).
//
// Here the 32 rays of a warp (camera rays of one tile, or the shadow rays of its hits towards one light) are walked
// TOGETHER:
//   * the node is fetched once per warp (5 x 128-bit loads at a warp-uniform address = one L1 wavefront each) and lane l
//     tests child (l & 7) against the BEAM of ray group (l >> 3), the 8 rays of one tile row: interval arithmetic on the
//     slab test with the group's origin box and inverse-direction box (mirrored so that every direction component is
//     positive). A child is entered when any group's beam may touch it; one redux.sync.or merges the 32 verdicts into the
//     hit word of bvh_query's traversal layout. 8 children x 4 beams in one SIMD pass instead of 8 children x 32 rays.
//   * the traversal stack is ONE stack per warp and lives in registers: entry i is held by lane i, pushed with a
//     predicated move and popped with two shuffles (the tree has at most 32 levels, rt_bvh_build.cu finish_build).
//   * triangles of a leaf are loaded once per warp; every lane filters them for its own ray (tri_filter, exact per ray),
//     collects the survivors in a bit mask and the reference-exact candidate test (exact_candidate) then runs on all
//     lanes' candidates together instead of per triangle with the few lanes that passed it.
// The beam test is conservative like every test of the filter stage (rt_bvh.cuh header): it may enter a child no ray
// hits, never skip one a ray hits; what is accepted is decided by exact_candidate with the reference's arithmetic, so
// ids, t and frames stay bit-identical to bvh_query's and the oracle's. Warps whose active rays do not share a
// direction octant (the tile over the image centre, rays after a bounce) fall back to bvh_query per lane.
#ifndef RTB200_BEAM_CUH
#define RTB200_BEAM_CUH

#include "rt_bvh.cuh"

namespace rtd {

#ifdef __CUDACC__

// relative widening of every interval bound: covers the roundings of plane - origin, the product with 1 / d and of
// 1 / d itself (each <= 2^-23 relative), far below the 2e-4 world units every leaf box is inflated by
#define RT_BEAM_SLACK 1.9073486328125e-06f  /* 2^-19 */

struct BeamStats { unsigned node_fetches; };

__device__ __forceinline__ float beam_gmin(float v) {  // over the 8 lanes of a ray group
    v = fminf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    v = fminf(v, __shfl_xor_sync(0xffffffffu, v, 2));
    return fminf(v, __shfl_xor_sync(0xffffffffu, v, 4));
}
__device__ __forceinline__ float beam_gmax(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
    return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 4));
}

// The per-lane walk for warps whose rays do not share a direction octant: out of line, it is cold.
template <bool STATS, bool FAST>
__device__ __noinline__ bool beam_fallback(const SceneView& s, const rtx::Ray& r, bool any, bool use_box, float max_t, Hit& best, TravStats& st) {
    TravStats* stp = STATS ? &st : nullptr;
    return any ? bvh_query<true, FAST>(s, r, use_box, max_t, best, stp) : bvh_query<false, FAST>(s, r, use_box, max_t, best, stp);
}

// Closest-hit (any == false) or occlusion (any == true) query for the rays of one warp. Every lane of the warp must
// call it (full-mask shuffles); `active` says whether the lane carries a ray. Closest hit: returns whether the lane's
// ray hit, `best` holds the hit. Occlusion: returns whether anything blocks the lane's ray within max_t.
// One out-of-line copy serves the camera rays and every light's shadow rays of k_tile (`any` is a run-time flag): the
// kernel's hot loop has to fit the instruction cache (the first version inlined a closest-hit and an occlusion copy, each
// with its per-lane fallback — 96 KB of SASS, 3.2 warps stalled on instruction fetch per issued instruction).
template <bool STATS, bool FAST = false>
__device__ __noinline__ bool beam_query(const SceneView& s, const rtx::Ray& r, bool active, const bool ANY, bool use_box, float max_t,
                                        Hit& best, TravStats& st, BeamStats& bst) {
    const unsigned FULLM = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    hit_reset(best);
    if (s.n_prims <= 0) return false;
    const unsigned act0 = __ballot_sync(FULLM, active);
    if (act0 == 0u) return false;
    RayPrep rp = ray_prep(r);
    // one direction octant for all active rays? (octinv of the first active lane)
    const unsigned oct_first = __shfl_sync(FULLM, rp.octinv, __ffs((int)act0) - 1);
    if (__any_sync(FULLM, active && rp.octinv != oct_first)) {
        if (!active) return false;
        return beam_fallback<STATS, FAST>(s, r, ANY, use_box, max_t, best, st);
    }
    const unsigned octinv = oct_first;
    // mirrored coordinates: axis k negated when the rays travel in its negative direction, so that 1 / d > 0
    const bool nx = !(octinv & 1u), ny = !(octinv & 2u), nz = !(octinv & 4u);  // octinv = 7 - octant
    const float INF = INFINITY;
    const float mox = nx ? -rp.o.x : rp.o.x, moy = ny ? -rp.o.y : rp.o.y, moz = nz ? -rp.o.z : rp.o.z;
    const float mix = fabsf(rp.idir.x), miy = fabsf(rp.idir.y), miz = fabsf(rp.idir.z);
    // beam of this lane's ray group (the 8 lanes of one tile row): origin box and inverse-direction box
    const float Olx = beam_gmin(active ? mox : INF), Ohx = beam_gmax(active ? mox : -INF);
    const float Oly = beam_gmin(active ? moy : INF), Ohy = beam_gmax(active ? moy : -INF);
    const float Olz = beam_gmin(active ? moz : INF), Ohz = beam_gmax(active ? moz : -INF);
    const float Ilx = beam_gmin(active ? mix : INF), Ihx = beam_gmax(active ? mix : -INF);
    const float Ily = beam_gmin(active ? miy : INF), Ihy = beam_gmax(active ? miy : -INF);
    const float Ilz = beam_gmin(active ? miz : INF), Ihz = beam_gmax(active ? miz : -INF);
    const bool group_has_rays = beam_gmax(active ? 1.0f : 0.0f) > 0.0f;
    float tcull = ANY ? fmaf(max_t, RT_TCULL_REL, RT_TCULL_ABS) : INF;  // this lane's ray
    float tcull_g = beam_gmax(active ? tcull : -INF);                    // its group's beam
    const int child = lane & 7;
    InstCtx ctx;
    ctx.inst = -1;
    bool occluded = false;
    uint2 my_entry = make_uint2(0u, 0u);  // stack entry `lane` of the warp's stack
    int sp = 0;
    uint2 ng;  // node group: x = base index, y = hit bits 24..31 | imask
    ng.x = 0u; ng.y = 0x80000000u;
    for (;;) {
        uint2 tg;  // triangle group: x = prim base, y = 24 hit bits
        if (ng.y > 0x00ffffffu) {
            unsigned hits = ng.y;
            const unsigned bit = 31u - (unsigned)__clz((int)hits);
            hits &= ~(1u << bit);
            if (hits & 0xff000000u) {
                if (lane == sp) { my_entry.x = ng.x; my_entry.y = hits; }
                sp++;
            }
            const unsigned slot = (bit - 24u) ^ octinv;
            const unsigned rel = (unsigned)__popc(hits & 0xffu & ((1u << slot) - 1u));
            const uint4* np = s.nodes + (size_t)(ng.x + rel) * 5;
            const uint4 q0 = ldg4u(np), q1 = ldg4u(np + 1), q2 = ldg4u(np + 2), q3 = ldg4u(np + 3), q4 = ldg4u(np + 4);
            if (STATS) { bst.node_fetches += lane == 0 ? 1u : 0u; st.nodes += active ? 1u : 0u; }
            // this lane's child: meta byte and the six quantised planes
            const unsigned meta = __byte_perm(q1.z, q1.w, child) & 0xffu;
            const float px = __uint_as_float(q0.x), py = __uint_as_float(q0.y), pz = __uint_as_float(q0.z);
            const float sx = __uint_as_float((q0.w & 0xffu) << 23), sy = __uint_as_float(((q0.w >> 8) & 0xffu) << 23),
                        sz = __uint_as_float(((q0.w >> 16) & 0xffu) << 23);
            const float lox = fmaf((float)(__byte_perm(q2.x, q2.y, child) & 0xffu), sx, px);
            const float loy = fmaf((float)(__byte_perm(q2.z, q2.w, child) & 0xffu), sy, py);
            const float loz = fmaf((float)(__byte_perm(q3.x, q3.y, child) & 0xffu), sz, pz);
            const float hix = fmaf((float)(__byte_perm(q3.z, q3.w, child) & 0xffu), sx, px);
            const float hiy = fmaf((float)(__byte_perm(q4.x, q4.y, child) & 0xffu), sy, py);
            const float hiz = fmaf((float)(__byte_perm(q4.z, q4.w, child) & 0xffu), sz, pz);
            const float nearx = nx ? -hix : lox, farx = nx ? -lox : hix;
            const float neary = ny ? -hiy : loy, fary = ny ? -loy : hiy;
            const float nearz = nz ? -hiz : loz, farz = nz ? -loz : hiz;
            // entry: the smallest (plane - origin) x the matching end of the 1 / d interval, widened downwards
            const float ax = nearx - Ohx, ay = neary - Ohy, az = nearz - Ohz;
            float t0x = ax * (ax >= 0.0f ? Ilx : Ihx), t0y = ay * (ay >= 0.0f ? Ily : Ihy), t0z = az * (az >= 0.0f ? Ilz : Ihz);
            t0x = fmaf(-fabsf(t0x), RT_BEAM_SLACK, t0x); t0y = fmaf(-fabsf(t0y), RT_BEAM_SLACK, t0y); t0z = fmaf(-fabsf(t0z), RT_BEAM_SLACK, t0z);
            // exit: the largest (plane - origin), widened upwards
            const float bx = farx - Olx, by = fary - Oly, bz = farz - Olz;
            float t1x = bx * (bx >= 0.0f ? Ihx : Ilx), t1y = by * (by >= 0.0f ? Ihy : Ily), t1z = bz * (bz >= 0.0f ? Ihz : Ilz);
            t1x = fmaf(fabsf(t1x), RT_BEAM_SLACK, t1x); t1y = fmaf(fabsf(t1y), RT_BEAM_SLACK, t1y); t1z = fmaf(fabsf(t1z), RT_BEAM_SLACK, t1z);
            const float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
            const float tmax = fminf(fminf(t1x, t1y), fminf(t1z, tcull_g));
            // traversal layout of this child's bits: inner children at 24 + (slot ^ octinv), leaves at their triangle offset
            const unsigned is_inner = (meta & (meta << 1)) & 0x10u;
            const unsigned bit_index = (meta ^ (is_inner ? octinv : 0u)) & 0x1fu;
            const unsigned bits = ((meta >> 5) & 7u) << bit_index;
            const bool enter = group_has_rays && meta != 0u && tmin <= tmax;
            const unsigned h = __reduce_or_sync(FULLM, enter ? bits : 0u);
            ng.x = q1.x; ng.y = (h & 0xff000000u) | (q0.w >> 24);
            tg.x = q1.y; tg.y = h & 0x00ffffffu;
        } else {
            tg = ng;
            ng.x = 0u; ng.y = 0u;
        }
        if (tg.y) {
            // every lane filters the leaf's triangles for its own ray; survivors wait in `cand`
            unsigned cand = 0u;
            unsigned todo = tg.y;
            bool fast_changed = false;
            while (todo) {
                const unsigned k = (unsigned)__ffs((int)todo) - 1u;
                todo &= todo - 1u;
                const float4* pp = s.prims + (size_t)(tg.x + k) * 3;
                const float4 p0 = ldg4f(pp), p1 = ldg4f(pp + 1), p2 = ldg4f(pp + 2);
                if (active) {
                    if (STATS) st.prims++;
                    if (FAST) {
                        float ft, fu, fv;
                        const int verdict = tri_filter_clear(rp.o, rp.d, p0, p1, p2, tcull, ft, fu, fv);
                        if (verdict == 2 && (!ANY || ft <= max_t * 0.9999f)) {
                            // tolerance mode: a clear hit needs no reference-exact test
                            if (fast_candidate(__float_as_int(p0.w), __float_as_int(p1.w), ft, fu, fv, best)) {
                                if (ANY) { occluded = true; active = false; cand = 0u; tcull = -INF; fast_changed = true; }
                                else { tcull = fmaf(best.t, RT_TCULL_REL, RT_TCULL_ABS); fast_changed = true; }
                            }
                        } else if (verdict) {
                            cand |= 1u << k;
                        }
                    } else if (tri_filter(rp.o, rp.d, p0, p1, p2, tcull)) {
                        cand |= 1u << k;
                    }
                }
            }
            // the reference-exact test, one candidate per lane and round, all lanes that have one together
            bool changed = fast_changed;
            while (__any_sync(FULLM, cand != 0u)) {
                if (cand) {
                    const unsigned k = (unsigned)__ffs((int)cand) - 1u;
                    cand &= cand - 1u;
                    const float4* pp = s.prims + (size_t)(tg.x + k) * 3;
                    const int inst = __float_as_int(ldg4f(pp).w), tri = __float_as_int(ldg4f(pp + 1).w);
                    if (STATS) st.exact++;
                    if (exact_candidate(s, r, inst, tri, use_box, ctx, best)) {
                        if (ANY) {
                            if (!(best.t > max_t)) { occluded = true; active = false; cand = 0u; tcull = -INF; changed = true; }
                            else hit_reset(best);  // accepted but beyond the light: keep looking
                        } else {
                            tcull = fmaf(best.t, RT_TCULL_REL, RT_TCULL_ABS);
                            changed = true;
                        }
                    }
                }
            }
            if (__any_sync(FULLM, changed)) {
                if (ANY && !__any_sync(FULLM, active)) break;  // every ray of the warp is answered
                tcull_g = beam_gmax(active ? tcull : -INF);
            }
        }
        if (ng.y <= 0x00ffffffu) {
            if (sp == 0) break;
            sp--;
            ng.x = __shfl_sync(FULLM, my_entry.x, sp);
            ng.y = __shfl_sync(FULLM, my_entry.y, sp);
        }
    }
    return ANY ? occluded : best.inst >= 0;
}

#endif  // __CUDACC__

}  // namespace rtd
#endif
