// rt_bvh_build.cu — GPU LBVH build and collapse into the compressed BVH8.
//
// Replaces rtracer::gpu::create_bvh / create_boxes (src/raytracer.cu:54-89) and
// ropt::gpu::BVH::BVH: gen_morton + gen_numbers + thrust::sort_by_key + reorder + the per-level
// build_bvh_layer launches (src/rayopt/bvh.cu:11-91). Pipeline (all on one stream, no host
// round trips, no allocation after the first frame):
//
//   k_flatten      world-space triangles + inflated AABBs + scene bounds      (create_boxes)
//   k_morton       63-bit quantised Morton code of each centroid, iota        (gen_morton, gen_numbers)
//   radix sort     8 x (k_radix_hist, k_radix_scan, k_radix_scatter), LSD, stable (thrust::sort_by_key)
//   k_karras       binary radix tree topology, Karras 2012                    (build_bvh_layer)
//   k_refit        bottom-up AABB union with arrival counters                 (merge_into)
//   k_collapse     level-by-level BVH2 -> BVH8, quantisation, slot assignment, leaf-ordered triangles
//
// Every step is deterministic (stable sort, exact min/max, fixed float evaluation order with
// -fmad=false), so the result is bit-identical to the sequential host build in oracle/lbvh_host.c
// that tests/test_gpu_parity.py::test_bvh_build_bit_exact_vs_host compares it against.
#include <cooperative_groups.h>
#include <cstdlib>
#include <cstring>
#include <string>

#include "rt_bvh.cuh"
#include "rt_internal.h"

namespace cg = cooperative_groups;
using namespace rtd;

namespace {

constexpr int kThreads = 256;

// ---------------------------------------------------------------- flatten
__device__ __forceinline__ int float_flip(float f) {  // order-preserving float -> int
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float float_unflip(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void k_bounds_init(int* bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0x7fffffff;
    else if (threadIdx.x < 6) bounds[threadIdx.x] = (int)0x80000000;
}

// World-space triangle p = prim_offset[inst] + k: record, inflated box, tight bounds (lo / hi).
__device__ __forceinline__ void flatten_prim(const SceneView& s, const int* prim_offset, int p, float4* prim_world, float* prim_box,
                                             float lo[3], float hi[3]) {
    {
        int a = 0, b = s.n_insts;  // last instance with prim_offset[inst] <= p
        while (b - a > 1) {
            int m = (a + b) >> 1;
            if (prim_offset[m] <= p) a = m; else b = m;
        }
        int inst = a;
        const InstRec& I = s.insts[inst];
        const MeshRec& M = s.meshes[I.mesh];
        int tri = M.tri_begin + (p - prim_offset[inst]);
        float L;
        if (M.kind == RT_HITABLE_SPHERE) {
            // analytic sphere: world centre, zero edges and L = +inf, which tri_filter passes unconditionally
            // (the leaf box is the filter; the exact sphere test decides). Tight box = centre +- radius.
            const float3 c = rtx::point_from_local(I.pose, rtx::point_from_local(M.pose, rtx::mk3(0.f, 0.f, 0.f)));
            prim_world[3 * (size_t)p + 0] = make_float4(c.x, c.y, c.z, __int_as_float(inst));
            prim_world[3 * (size_t)p + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(tri));
            prim_world[3 * (size_t)p + 2] = make_float4(0.f, 0.f, 0.f, INFINITY);
            lo[0] = c.x - M.radius; lo[1] = c.y - M.radius; lo[2] = c.z - M.radius;
            hi[0] = c.x + M.radius; hi[1] = c.y + M.radius; hi[2] = c.z + M.radius;
            L = 6.0f * M.radius;
        } else {
            const TriRec T = s.tris[tri];
            float3 w[3];
            const int vid[3] = {T.v0, T.v1, T.v2};
#pragma unroll
            for (int k = 0; k < 3; k++) {
                float4 v = s.verts[vid[k]];
                w[k] = rtx::point_from_local(I.pose, rtx::point_from_local(M.pose, rtx::mk3(v.x, v.y, v.z)));
            }
            float3 e1 = rtx::sub3(w[1], w[0]), e2 = rtx::sub3(w[2], w[0]);
            L = ((fabsf(e1.x) + fabsf(e1.y)) + fabsf(e1.z)) + ((fabsf(e2.x) + fabsf(e2.y)) + fabsf(e2.z));
            prim_world[3 * (size_t)p + 0] = make_float4(w[0].x, w[0].y, w[0].z, __int_as_float(inst));
            prim_world[3 * (size_t)p + 1] = make_float4(e1.x, e1.y, e1.z, __int_as_float(tri));
            prim_world[3 * (size_t)p + 2] = make_float4(e2.x, e2.y, e2.z, L);
            lo[0] = fminf(fminf(w[0].x, w[1].x), w[2].x); hi[0] = fmaxf(fmaxf(w[0].x, w[1].x), w[2].x);
            lo[1] = fminf(fminf(w[0].y, w[1].y), w[2].y); hi[1] = fmaxf(fmaxf(w[0].y, w[1].y), w[2].y);
            lo[2] = fminf(fminf(w[0].z, w[1].z), w[2].z); hi[2] = fmaxf(fmaxf(w[0].z, w[1].z), w[2].z);
        }
        // stored box = tight box inflated by the acceptance margin of tri_filter plus coordinate rounding
        float amax = fmaxf(fmaxf(fmaxf(fabsf(lo[0]), fabsf(hi[0])), fmaxf(fabsf(lo[1]), fabsf(hi[1]))),
                           fmaxf(fabsf(lo[2]), fabsf(hi[2])));
        float m = (2e-4f + 4e-5f * L) + 1e-6f * amax;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            prim_box[6 * (size_t)p + k] = lo[k] - m;
            prim_box[6 * (size_t)p + 3 + k] = hi[k] + m;
        }
    }
}

// One thread per world-space triangle.
__global__ void k_flatten(SceneView s, const int* __restrict__ prim_offset, int n_prims, float4* __restrict__ prim_world,
                          float* __restrict__ prim_box, int* __restrict__ bounds) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (p < n_prims) flatten_prim(s, prim_offset, p, prim_world, prim_box, lo, hi);
    // block reduce of the TIGHT bounds, then one atomic per block and component
    __shared__ int s_lo[3][kThreads / 32], s_hi[3][kThreads / 32];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        int l = float_flip(lo[k]), h = float_flip(hi[k]);
        for (int o = 16; o; o >>= 1) {
            l = min(l, __shfl_xor_sync(0xffffffffu, l, o));
            h = max(h, __shfl_xor_sync(0xffffffffu, h, o));
        }
        if (lane == 0) { s_lo[k][warp] = l; s_hi[k][warp] = h; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        int k = threadIdx.x % 3;
        bool is_hi = threadIdx.x >= 3;
        int v = is_hi ? s_hi[k][0] : s_lo[k][0];
        for (int w2 = 1; w2 < kThreads / 32; w2++) v = is_hi ? max(v, s_hi[k][w2]) : min(v, s_lo[k][w2]);
        if (is_hi) atomicMax(&bounds[3 + k], v); else atomicMin(&bounds[k], v);
    }
}

// ---------------------------------------------------------------- Morton codes
__device__ __forceinline__ unsigned long long expand21(unsigned long long v) {
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x001f00000000ffffull;
    v = (v | v << 16) & 0x001f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

// 63-bit Morton code of triangle p's centroid inside the scene bounds.
// key_shift: 0 keeps the 63-bit code; 31 keeps its top 32 bits (scenes up to 2^20 triangles: four
// radix passes instead of eight; equal prefixes fall back to the triangle index like equal codes do).
__device__ __forceinline__ unsigned long long morton_key(const float4* prim_world, int p, const float blo[3], const float bhi[3],
                                                         int key_shift) {
    // centroid of the tight box, recomputed from the triangle so it does not depend on the margin
    float4 p0 = prim_world[3 * (size_t)p], p1 = prim_world[3 * (size_t)p + 1], p2 = prim_world[3 * (size_t)p + 2];
    float v0[3] = {p0.x, p0.y, p0.z}, e1[3] = {p1.x, p1.y, p1.z}, e2[3] = {p2.x, p2.y, p2.z};
    unsigned long long q[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float a = v0[k], b = v0[k] + e1[k], c = v0[k] + e2[k];
        float lo = fminf(fminf(a, b), c), hi = fmaxf(fmaxf(a, b), c);
        float cen = (lo + hi) * 0.5f;
        float ext = bhi[k] - blo[k];
        float n = ext > 0.0f ? (cen - blo[k]) / ext : 0.0f;
        float sc = n * 2097152.0f;
        sc = fminf(fmaxf(sc, 0.0f), 2097151.0f);
        q[k] = (unsigned long long)(unsigned)sc;
    }
    return ((expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2])) >> key_shift;
}

__global__ void k_morton(const float* __restrict__ prim_box, const float4* __restrict__ prim_world, int n_prims,
                         const int* __restrict__ bounds, unsigned long long* __restrict__ keys, int* __restrict__ vals,
                         float* __restrict__ bounds_f, int key_shift) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    float blo[3], bhi[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { blo[k] = float_unflip(bounds[k]); bhi[k] = float_unflip(bounds[3 + k]); }
    if (p == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) { bounds_f[k] = blo[k]; bounds_f[3 + k] = bhi[k]; }
    }
    if (p >= n_prims) return;
    keys[p] = morton_key(prim_world, p, blo, bhi, key_shift);
    vals[p] = p;
    (void)prim_box;
}

// ---------------------------------------------------------------- LSD radix sort (8 bits/pass, stable)
constexpr int kRadixTile = 2048;  // keys per block: 256 threads, 8 warps x 256 consecutive keys
constexpr int kBins = 256;

__global__ void k_radix_hist(const unsigned long long* __restrict__ keys, int n, int shift, unsigned* __restrict__ hist,
                             int n_blocks) {
    __shared__ unsigned sh[kBins];
    sh[threadIdx.x] = 0;
    __syncthreads();
    int base = blockIdx.x * kRadixTile;
    for (int i = threadIdx.x; i < kRadixTile; i += kThreads) {
        int idx = base + i;
        if (idx < n) atomicAdd(&sh[(unsigned)(keys[idx] >> shift) & 0xffu], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * n_blocks + blockIdx.x] = sh[threadIdx.x];  // bin-major for the scan
}

// Many tiles: block d turns digit d's per-tile counts (hist[d * n_blocks ..]) into exclusive prefixes over the tiles
// and leaves the digit's total in hist[kBins * n_blocks + d]; k_radix_scatter adds the prefix over the digit totals.
// (One block scanning all kBins * n_blocks counters took 1.4 ms per pass at 3.8 M keys: 55 % of that frame.)
__global__ void __launch_bounds__(kThreads) k_radix_scan(unsigned* __restrict__ hist, int n_blocks) {
    __shared__ unsigned warp_sum[kThreads / 32];
    __shared__ unsigned carry;
    unsigned* col = hist + (size_t)blockIdx.x * n_blocks;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n_blocks; base += kThreads) {
        int idx = base + threadIdx.x;
        unsigned v = idx < n_blocks ? col[idx] : 0u;
        unsigned x = v;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sum[warp] = x;
        __syncthreads();
        unsigned woff = 0;
        for (int w = 0; w < warp; w++) woff += warp_sum[w];
        unsigned c = carry;
        if (idx < n_blocks) col[idx] = c + woff + x - v;
        __syncthreads();
        if (threadIdx.x == kThreads - 1) carry = c + woff + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) hist[(size_t)kBins * n_blocks + blockIdx.x] = carry;
}

// FUSED_SCAN: `hist` holds raw per-block counts and every block derives its own global offsets
// (few blocks: cheaper than a separate scan launch); otherwise k_radix_scan left per-digit prefixes over the
// tiles plus the digit totals, and the block only scans the 256 totals.
template <bool FUSED_SCAN>
__global__ void k_radix_scatter(const unsigned long long* __restrict__ keys_in, const int* __restrict__ vals_in, int n,
                                int shift, const unsigned* __restrict__ hist, int n_blocks,
                                unsigned long long* __restrict__ keys_out, int* __restrict__ vals_out) {
    __shared__ unsigned wcount[kThreads / 32][kBins];  // per-warp digit counts, then running offsets
    __shared__ unsigned s_wtot[kThreads / 32];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (kThreads / 32) * kBins; i += kThreads) (&wcount[0][0])[i] = 0;
    __syncthreads();
    int wbase = blockIdx.x * kRadixTile + warp * (kRadixTile / (kThreads / 32));
    constexpr int kIters = kRadixTile / (kThreads / 32) / 32;
    unsigned long long k[kIters];
    int v[kIters];
#pragma unroll
    for (int it = 0; it < kIters; it++) {
        int idx = wbase + it * 32 + lane;
        bool ok = idx < n;
        k[it] = ok ? keys_in[idx] : ~0ull;
        v[it] = ok ? vals_in[idx] : 0;
        unsigned d = (unsigned)(k[it] >> shift) & 0xffu;
        unsigned act = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            unsigned peers = __match_any_sync(act, d);
            if (lane == (31 - __clz((int)peers))) wcount[warp][d] += (unsigned)__popc(peers);
        }
        __syncwarp();
    }
    __syncthreads();
    unsigned gbase;
    {   // exclusive scan over (digit, block) in bin-major order, evaluated at (d, blockIdx.x)
        unsigned d = threadIdx.x, tot = 0, before = 0;
        if (FUSED_SCAN) {
            for (int b2 = 0; b2 < n_blocks; b2++) {
                unsigned c = hist[(size_t)d * n_blocks + b2];
                tot += c;
                if (b2 < (int)blockIdx.x) before += c;
            }
        } else {
            tot = hist[(size_t)kBins * n_blocks + d];
            before = hist[(size_t)d * n_blocks + blockIdx.x];
        }
        unsigned x = tot;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_wtot[warp] = x;
        __syncthreads();
        unsigned woff = 0;
        for (int w = 0; w < warp; w++) woff += s_wtot[w];
        gbase = woff + x - tot + before;
    }
    {   // thread d: turn counts into block-local exclusive offsets across warps, plus the global base
        unsigned d = threadIdx.x;
        unsigned run = gbase;
        for (int w = 0; w < kThreads / 32; w++) {
            unsigned c = wcount[w][d];
            wcount[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < kIters; it++) {
        int idx = wbase + it * 32 + lane;
        bool ok = idx < n;
        unsigned d = (unsigned)(k[it] >> shift) & 0xffu;
        unsigned act = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            unsigned peers = __match_any_sync(act, d);
            unsigned rank = (unsigned)__popc(peers & ((1u << lane) - 1u));
            unsigned pos = wcount[warp][d] + rank;
            keys_out[pos] = k[it];
            vals_out[pos] = v[it];
            __syncwarp(act);
            if (lane == (31 - __clz((int)peers))) wcount[warp][d] += (unsigned)__popc(peers);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- Karras 2012 hierarchy
// Node ids: internal 0..n-2 (root 0), leaf of sorted position j = n-1+j.
__device__ __forceinline__ int delta(const unsigned long long* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}
// The same function on the low words of keys that are known to be below 2^32 (32-bit Morton prefixes, key_shift 31):
// clzll(a ^ b) = 32 + clz(lo(a) ^ lo(b)). Lets the one-launch build keep the sorted keys in shared memory.
__device__ __forceinline__ int delta(const unsigned* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return 32 + __clz((int)(a ^ b));
}

template <class K>
__device__ __forceinline__ void karras_node(const K* keys, int n, int i, int* left, int* right, int* parent,
                                            int* range_lo, int* range_hi) {
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) > 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int lc = (lo == gamma) ? (n - 1 + gamma) : gamma;
    int rc = (hi == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
    left[i] = lc;
    right[i] = rc;
    parent[lc] = i;
    parent[rc] = i;
    range_lo[i] = lo;
    range_hi[i] = hi;
    if (i == 0) parent[0] = -1;
}

__global__ void k_karras(const unsigned long long* __restrict__ keys, int n, int* __restrict__ left, int* __restrict__ right,
                         int* __restrict__ parent, int* __restrict__ range_lo, int* __restrict__ range_hi) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    karras_node(keys, n, i, left, right, parent, range_lo, range_hi);
}

// One thread per leaf climbs; the second arrival at a node unions its children.
__global__ void k_refit(const int* __restrict__ sorted_prim, const float* __restrict__ prim_box, int n,
                        const int* __restrict__ left, const int* __restrict__ right, const int* __restrict__ parent,
                        unsigned* __restrict__ arrive, float* __restrict__ node_box) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int p = sorted_prim[j];
    int node = n - 1 + j;
#pragma unroll
    for (int k = 0; k < 6; k++) node_box[6 * (size_t)node + k] = prim_box[6 * (size_t)p + k];
    if (n == 1) return;
    __threadfence();
    int cur = parent[node];
    while (cur >= 0) {
        if (atomicAdd(&arrive[cur], 1u) == 0u) return;
        __threadfence();
        int l = left[cur], r = right[cur];
        volatile const float* bl = node_box + 6 * (size_t)l;
        volatile const float* br = node_box + 6 * (size_t)r;
        float o[6];
#pragma unroll
        for (int k = 0; k < 3; k++) { o[k] = fminf(bl[k], br[k]); o[3 + k] = fmaxf(bl[3 + k], br[3 + k]); }
#pragma unroll
        for (int k = 0; k < 6; k++) node_box[6 * (size_t)cur + k] = o[k];
        __threadfence();
        cur = parent[cur];
    }
}

// ---------------------------------------------------------------- BVH2 -> BVH8 collapse
struct CollapseArgs {
    int n;                 // primitives
    const int* left; const int* right; const int* range_lo; const int* range_hi;
    const float* node_box; // 6 floats per BVH2 node
    const int* sorted_prim;
    const float4* prim_world;
    uint4* nodes8;         // out
    float4* prims8;        // out
    int* prim_order8;      // out: sorted position of each BVH8 triangle slot
    int* level_a; int* level_b;  // BVH2 ids of the nodes of the current / next level
    int* slots;            // 8 ints per node of the current level: BVH2 child id per slot or -1
    int* counts;           // 2 ints per node of the current level: inner children, leaf triangles
    int* block_tot;        // [2 parities][gridDim.x][2]
    int* result;           // [0] = n_nodes8, [1] = n_prims emitted, [2] = levels, [3] = max level width
    int* node8_root;       // out: BVH2 id each BVH8 node was collapsed from      (kept for k_refit8)
    int* node8_slot;       // out: BVH2 id in each of its 8 slots, -1 when empty
    const uint4* rec;      // per-node records (collapse_children_rec) or null: read the separate arrays
    unsigned long long* lvl_ns;  // RT_BUILD_TIMERS: per level, time stamps of block 0 after phase A and after phase B; else null
};

__device__ __forceinline__ int node_size(const CollapseArgs& a, int id) {
    return id >= a.n - 1 ? 1 : a.range_hi[id] - a.range_lo[id] + 1;
}
__device__ __forceinline__ int node_first(const CollapseArgs& a, int id) { return id >= a.n - 1 ? id - (a.n - 1) : a.range_lo[id]; }
__device__ __forceinline__ float half_area(const float* b) {
    float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
    return (dx * dy + dy * dz) + dz * dx;
}

// ---- 8 lanes per BVH8 node (4 nodes per warp). Every shuffle below is executed by all 32 lanes
// with uniform trip counts; `width = 8` keeps the exchange inside one node's lane group. ----
#define FULL 0xffffffffu
__device__ __forceinline__ int g_shfl(int v, int src) { return __shfl_sync(FULL, v, src, 8); }
__device__ __forceinline__ int g_sum(int v) {
    v += __shfl_xor_sync(FULL, v, 1, 8);
    v += __shfl_xor_sync(FULL, v, 2, 8);
    v += __shfl_xor_sync(FULL, v, 4, 8);
    return v;
}
__device__ __forceinline__ unsigned g_or(unsigned v) {
    v |= __shfl_xor_sync(FULL, v, 1, 8);
    v |= __shfl_xor_sync(FULL, v, 2, 8);
    v |= __shfl_xor_sync(FULL, v, 4, 8);
    return v;
}
// exclusive prefix sum over the 8 lanes of a group
__device__ __forceinline__ int g_excl(int v, int j) {
    int x = v;
    for (int o = 1; o < 8; o <<= 1) {
        int y = __shfl_up_sync(FULL, x, o, 8);
        if (j >= o) x += y;
    }
    return x - v;
}
// argmax of (val, lowest lane on ties) over the group; returns the winning lane, val_out its value
__device__ __forceinline__ int g_argmax(float val, int j, float& val_out) {
    int idx = j;
    for (int o = 1; o < 8; o <<= 1) {
        float ov = __shfl_xor_sync(FULL, val, o, 8);
        int oi = __shfl_xor_sync(FULL, idx, o, 8);
        if (ov > val || (ov == val && oi < idx)) { val = ov; idx = oi; }
    }
    val_out = val;
    return idx;
}

// Children of one BVH8 node and their slots, computed by the 8 lanes of a group (lane j holds list
// position j). Rule (identical to the sequential host build): start from the two BVH2 children;
// while fewer than 8, replace the child of largest half-area among those covering more than 3
// triangles by its own two children (left in place, right appended); ties go to the earlier child.
// Slots: greedy assignment maximising (child centroid - node centroid) . octant direction of the
// slot, scanning children then slots with a strict comparison. On return lane s holds the BVH2 id
// in slot s (or -1), and n_inner / n_leaf_prims are group-uniform.
// Greedy slot assignment over the 8 lanes of a group, one (child, slot) pair per round: lane j holds child j's cost row
// (has_child: the lane holds a child); returns the slot of this lane's child or -1.
__device__ __forceinline__ int greedy_slots_g(const float (&cost)[8], bool has_child, int j) {
    unsigned slot_done = 0;
    bool mine_done = !has_child;
    int my_slot = -1;
#pragma unroll 1
    for (int it = 0; it < 8; it++) {
        // this lane's best free slot (lowest slot on ties)
        float bv = -INFINITY;
        int bs = -1;
#pragma unroll
        for (int s = 0; s < 8; s++)
            if (!(slot_done & (1u << s)) && (bs < 0 || cost[s] > bv)) { bv = cost[s]; bs = s; }
        // lanes that are done or empty must never win: key below every real value, NaN-free
        float key = mine_done ? -INFINITY : bv;
        int cand = mine_done ? 8 + j : j;  // done lanes sort after live ones on ties at -inf
        // argmax over lanes with ties to the lowest live lane
        float kv = key;
        int ki = cand;
        for (int o = 1; o < 8; o <<= 1) {
            float ov = __shfl_xor_sync(FULL, kv, o, 8);
            int oi = __shfl_xor_sync(FULL, ki, o, 8);
            if (ov > kv || (ov == kv && oi < ki)) { kv = ov; ki = oi; }
        }
        bool any = ki < 8;  // a live lane won
        int wslot = g_shfl(bs, ki & 7);
        if (any) {
            if (j == ki) { my_slot = bs; mine_done = true; }
            slot_done |= 1u << wslot;
        }
    }
    return my_slot;
}

// ---- the same selection from per-node RECORDS (the one-launch build): 64 bytes per BVH2 node,
//   q0 = box.lo.xyz, left      q1 = box.hi.xyz, right      q2 = lo, hi, area(left), area(right)      q3 = size(left), size(right), -, -
// with area(c) = half-area of c's box, or -1 when c covers <= 3 triangles (cannot be opened); leaves carry left = right = -1.
// One 64-byte load tells everything about a node AND whether its children can be opened, so an opening costs one round
// trip to L2 (the two new children fetch their records in the same branch) instead of four or five: the child lists of a
// level are chains of dependent loads, and those chains — not arithmetic — set the collapse's time at these scene sizes.
struct Rec2 { uint4 q0, q1, q2, q3; };
__device__ __forceinline__ void rec_store(uint4* rec, int id, const float* box, int left, int right, int lo, int hi, float al, float ar,
                                          int sl, int sr) {
    uint4* r = rec + 4 * (size_t)id;
    r[0] = make_uint4(__float_as_uint(box[0]), __float_as_uint(box[1]), __float_as_uint(box[2]), (unsigned)left);
    r[1] = make_uint4(__float_as_uint(box[3]), __float_as_uint(box[4]), __float_as_uint(box[5]), (unsigned)right);
    r[2] = make_uint4((unsigned)lo, (unsigned)hi, __float_as_uint(al), __float_as_uint(ar));
    r[3] = make_uint4((unsigned)sl, (unsigned)sr, 0u, 0u);
}
__device__ int collapse_children_rec(const CollapseArgs& a, int root, int j, int& n_inner, int& n_leaf_prims) {
    const uint4* rec = a.rec;
    int ch = -1, size = 0;
    float area = -2.0f;   // -2: empty position, -1: cannot be opened, >= 0: half-area
    // this lane's child: its box and what its record says about ITS children
    float cb[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int cl = -1, cr = -1, sl = 0, sr = 0;
    float al = -1.0f, ar = -1.0f;
    auto load = [&](int id) {
        const uint4* r = rec + 4 * (size_t)id;
        const uint4 q0 = r[0], q1 = r[1], q2 = r[2], q3 = r[3];
        cb[0] = __uint_as_float(q0.x); cb[1] = __uint_as_float(q0.y); cb[2] = __uint_as_float(q0.z);
        cb[3] = __uint_as_float(q1.x); cb[4] = __uint_as_float(q1.y); cb[5] = __uint_as_float(q1.z);
        cl = (int)q0.w; cr = (int)q1.w;
        al = __uint_as_float(q2.z); ar = __uint_as_float(q2.w);
        sl = (int)q3.x; sr = (int)q3.y;
    };
    // every lane reads the root's record (one broadcast transaction): its box for the cost rows, its children to start from
    float nb[6];
    int root_size, nc;
    {
        const uint4* r = rec + 4 * (size_t)root;
        const uint4 q0 = r[0], q1 = r[1], q2 = r[2], q3 = r[3];
        nb[0] = __uint_as_float(q0.x); nb[1] = __uint_as_float(q0.y); nb[2] = __uint_as_float(q0.z);
        nb[3] = __uint_as_float(q1.x); nb[4] = __uint_as_float(q1.y); nb[5] = __uint_as_float(q1.z);
        root_size = (int)q2.y - (int)q2.x + 1;
        const bool single = root >= a.n - 1 || root_size <= 3;  // whole scene is one leaf
        if (single) {
            if (j == 0) { ch = root; area = -1.0f; size = root >= a.n - 1 ? 1 : root_size; }
            nc = 1;
        } else {
            if (j == 0) { ch = (int)q0.w; area = __uint_as_float(q2.z); size = (int)q3.x; }
            if (j == 1) { ch = (int)q1.w; area = __uint_as_float(q2.w); size = (int)q3.y; }
            nc = 2;
        }
        if (ch >= 0) load(ch);
    }
#pragma unroll 1
    for (int it = 0; it < 6; it++) {
        float best_area;
        const int best = g_argmax(area, j, best_area);
        const bool open = best_area >= 0.0f && nc < 8;  // group-uniform; the trip count stays warp-uniform (full-mask shuffles)
        const int l = g_shfl(cl, best), r = g_shfl(cr, best);
        const float la = __int_as_float(g_shfl(__float_as_int(al), best)), ra = __int_as_float(g_shfl(__float_as_int(ar), best));
        const int ls = g_shfl(sl, best), rs = g_shfl(sr, best);
        if (open) {
            if (j == best) { ch = l; area = la; size = ls; }
            if (j == nc) { ch = r; area = ra; size = rs; }
            if (j == best || j == nc) load(ch);  // both new children fetch their records together: one round trip per opening
            nc++;
        }
    }
    // cost row of this lane's child
    float cost[8];
    {
        float ncx = (nb[0] + nb[3]) * 0.5f, ncy = (nb[1] + nb[4]) * 0.5f, ncz = (nb[2] + nb[5]) * 0.5f;
        float dx = 0.f, dy = 0.f, dz = 0.f;
        if (ch >= 0) {
            dx = (cb[0] + cb[3]) * 0.5f - ncx; dy = (cb[1] + cb[4]) * 0.5f - ncy; dz = (cb[2] + cb[5]) * 0.5f - ncz;
        }
#pragma unroll
        for (int s2 = 0; s2 < 8; s2++) cost[s2] = (((s2 & 1) ? dx : -dx) + ((s2 & 2) ? dy : -dy)) + ((s2 & 4) ? dz : -dz);
    }
    const int my_slot = greedy_slots_g(cost, ch >= 0, j);
    // lane s <- child assigned to slot s
    int child_of_slot = -1, sz = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        int sk = g_shfl(my_slot, k);
        int ck = g_shfl(ch, k);
        int zk = g_shfl(size, k);
        if (sk == j) { child_of_slot = ck; sz = zk; }
    }
    n_inner = g_sum(sz > 3 ? 1 : 0);
    n_leaf_prims = g_sum(sz > 3 ? 0 : sz);
    return child_of_slot;
}

__device__ int collapse_children_g(const CollapseArgs& a, int root, int j, int& n_inner, int& n_leaf_prims) {
    int ch = -1;          // BVH2 id at list position j
    int cl = -1, cr = -1; // its two children, fetched together with it (one memory round trip per opening)
    float area = -2.0f;   // -2: empty position, -1: cannot be opened, >= 0: half-area
    float cb[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int nc;
    auto load = [&](int id) {
        ch = id;
        const float* bx = a.node_box + 6 * (size_t)id;
#pragma unroll
        for (int k = 0; k < 6; k++) cb[k] = bx[k];
        if (id >= a.n - 1) {  // single triangle
            area = -1.0f;
        } else {
            int lo = a.range_lo[id], hi = a.range_hi[id];
            cl = a.left[id];
            cr = a.right[id];
            area = (hi - lo + 1) <= 3 ? -1.0f : half_area(cb);
        }
    };
    const bool single = root >= a.n - 1 || node_size(a, root) <= 3;  // whole scene is one leaf
    if (single) {
        if (j == 0) { load(root); area = -1.0f; }
        nc = 1;
    } else {
        if (j == 0) load(a.left[root]);
        if (j == 1) load(a.right[root]);
        nc = 2;
    }
#pragma unroll 1
    for (int it = 0; it < 6; it++) {
        float best_area;
        int best = g_argmax(area, j, best_area);
        int l = g_shfl(cl, best), r = g_shfl(cr, best);
        bool open = best_area >= 0.0f && nc < 8;  // group-uniform
        if (open) {
            if (j == best) load(l);
            if (j == nc) load(r);
            nc++;
        }
    }
    // cost row of this lane's child
    float cost[8];
    {
        const float* nb = a.node_box + 6 * (size_t)root;
        float ncx = (nb[0] + nb[3]) * 0.5f, ncy = (nb[1] + nb[4]) * 0.5f, ncz = (nb[2] + nb[5]) * 0.5f;
        float dx = 0.f, dy = 0.f, dz = 0.f;
        if (ch >= 0) {
            dx = (cb[0] + cb[3]) * 0.5f - ncx; dy = (cb[1] + cb[4]) * 0.5f - ncy; dz = (cb[2] + cb[5]) * 0.5f - ncz;
        }
#pragma unroll
        for (int s = 0; s < 8; s++) cost[s] = (((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy)) + ((s & 4) ? dz : -dz);
    }
    const int my_slot = greedy_slots_g(cost, ch >= 0, j);
    // lane s <- child assigned to slot s
    int child_of_slot = -1;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        int sk = g_shfl(my_slot, k);
        int ck = g_shfl(ch, k);
        if (sk == j) child_of_slot = ck;
    }
    int sz = child_of_slot >= 0 ? node_size(a, child_of_slot) : 0;
    n_inner = g_sum(sz > 3 ? 1 : 0);
    n_leaf_prims = g_sum(sz > 3 ? 0 : sz);
    return child_of_slot;
}

// Lane s of the group quantises and emits slot s of one BVH8 node; lanes 0 and 4 store the packed
// words. `c` = BVH2 id in this lane's slot or -1.
// BUILD == false (refit): topology, child_base and prim_base are the existing node's; only boxes and triangles change.
template <bool BUILD>
__device__ void write_node8_g(const CollapseArgs& a, int root, int node_idx, int c, int j, int child_base, int prim_base,
                              int* next_list, bool valid) {
    const float* nb = a.node_box + 6 * (size_t)root;
    float org[3] = {nb[0], nb[1], nb[2]};
    unsigned eb[3];
    float scale[3], inv_scale[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float ext = nb[3 + k] - nb[k];
        float f = ext / 255.0f;
        unsigned bits = __float_as_uint(f);
        unsigned e = (bits >> 23) & 0xffu;
        if (bits & 0x7fffffu) e += 1u;
        if (e < 1u) e = 1u;
        if (e > 253u) e = 253u;
        float sc = __uint_as_float(e << 23);
        if (org[k] + 255.0f * sc < nb[3 + k] && e < 253u) { e += 1u; sc = __uint_as_float(e << 23); }
        eb[k] = e;
        scale[k] = sc;
        inv_scale[k] = __uint_as_float((254u - e) << 23);
    }
    unsigned q[6] = {0, 0, 0, 0, 0, 0};
    int sz = 0;
    if (c >= 0) {
        const float* cb = a.node_box + 6 * (size_t)c;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float ql = floorf((cb[k] - org[k]) * inv_scale[k]);
            ql = fminf(fmaxf(ql, 0.0f), 255.0f);
            while (ql > 0.0f && org[k] + ql * scale[k] > cb[k]) ql -= 1.0f;
            float qh = ceilf((cb[3 + k] - org[k]) * inv_scale[k]);
            qh = fminf(fmaxf(qh, 0.0f), 255.0f);
            while (qh < 255.0f && org[k] + qh * scale[k] < cb[3 + k]) qh += 1.0f;
            q[k] = (unsigned)ql;
            q[3 + k] = (unsigned)qh;
        }
        sz = node_size(a, c);
    }
    const bool inner = sz > 3;
    const int inner_rank = g_excl(inner ? 1 : 0, j);
    const int prim_off = g_excl(inner ? 0 : sz, j);
    unsigned meta = 0;
    if (c >= 0) {
        if (inner) {
            meta = 0x20u | (24u + (unsigned)j);
            if (BUILD && valid) next_list[inner_rank] = c;
        } else {
            unsigned unary = sz == 1 ? 1u : (sz == 2 ? 3u : 7u);
            meta = (unary << 5) | (unsigned)prim_off;
            int first = node_first(a, c);
            for (int t = 0; t < sz; t++) {
                int dst = prim_base + prim_off + t;
                int p = a.sorted_prim[first + t];
                if (valid) {
                    if (BUILD) a.prim_order8[dst] = first + t;
                    a.prims8[3 * (size_t)dst + 0] = a.prim_world[3 * (size_t)p + 0];
                    a.prims8[3 * (size_t)dst + 1] = a.prim_world[3 * (size_t)p + 1];
                    a.prims8[3 * (size_t)dst + 2] = a.prim_world[3 * (size_t)p + 2];
                }
            }
        }
    }
    const unsigned imask = g_or(inner ? (1u << j) : 0u);
    // byte j of each 8-byte array: OR the shifted bytes of lanes 0..3 / 4..7 into one word each
    const int sh = 8 * (j & 3);
    auto pack = [&](unsigned v) {
        unsigned w = v << sh;
        w |= __shfl_xor_sync(FULL, w, 1, 8);
        w |= __shfl_xor_sync(FULL, w, 2, 8);
        return w;  // lanes 0..3 hold the low word, lanes 4..7 the high word
    };
    unsigned wm = pack(meta), w0 = pack(q[0]), w1 = pack(q[1]), w2 = pack(q[2]), w3 = pack(q[3]), w4 = pack(q[4]), w5 = pack(q[5]);
    if (!valid) return;
    if (BUILD) {
        a.node8_slot[8 * (size_t)node_idx + j] = c;
        if (j == 0) a.node8_root[node_idx] = root;
    }
    unsigned* out = (unsigned*)(a.nodes8 + 5 * (size_t)node_idx);
    if (j == 0) {
        out[0] = __float_as_uint(org[0]); out[1] = __float_as_uint(org[1]); out[2] = __float_as_uint(org[2]);
        out[3] = eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24);
        out[4] = (unsigned)child_base; out[5] = (unsigned)prim_base;
    }
    if (j == 0 || j == 4) {
        int h = j >> 2;  // 0: low words, 1: high words
        out[6 + h] = wm;
        out[8 + h] = w0; out[10 + h] = w1; out[12 + h] = w2;
        out[14 + h] = w3; out[16 + h] = w4; out[18 + h] = w5;
    }
}

// Grid barrier for the cooperative launch (all blocks co-resident): one arrival counter that only
// grows, so no reset race; cheaper than cooperative_groups' grid.sync() at these block counts.
__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned n_blocks, unsigned& epoch) {
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

// Cooperative kernel: all levels in one launch, two grid barriers per level. Blocks own
// contiguous chunks of the level so that the per-level exclusive scan (which fixes child_base
// and prim_base, i.e. the memory order of the whole tree) follows node order deterministically.
// Runs on G co-resident blocks (block index bid) that `sync` can synchronise: the whole grid of the cooperative
// k_collapse (global-atomic barrier) or the CTAs of one thread-block cluster inside k_build_cluster (hardware barrier).
struct GridSync {
    unsigned* bar;
    unsigned n_blocks, epoch;
    __device__ __forceinline__ void operator()() { grid_barrier(bar, n_blocks, epoch); }
};

__device__ __forceinline__ unsigned long long collapse_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

template <class Sync>
__device__ __forceinline__ void collapse_levels(const CollapseArgs& a, Sync& sync, const int G, const int bid) {
    __shared__ int s_warp[2][32];
    __shared__ int s_carry[2];
    __shared__ int s_base[2];
    const int B = blockDim.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int level_n = 1, level_base = 0, prim_total = 0, levels = 0, max_width = 1;
    if (bid == 0 && threadIdx.x == 0) a.level_a[0] = 0;
    sync();
    while (level_n > 0) {
        int parity = levels & 1;
        int* cur = parity ? a.level_b : a.level_a;
        int* next = parity ? a.level_a : a.level_b;
        int* btot = a.block_tot + (size_t)parity * G * 2;
        int per_block = (level_n + G - 1) / G;
        int c0 = min(level_n, bid * per_block), c1 = min(level_n, c0 + per_block);
#ifdef RT_FINE
#define FINE(k) if (a.lvl_ns && bid == 0 && threadIdx.x == 0 && levels == 2) a.lvl_ns[16 + (k)] = collapse_ns();
#else
#define FINE(k)
#endif
        FINE(0)
        // phase A: children + counts, 8 lanes per node
        const int group = threadIdx.x >> 3, j = threadIdx.x & 7, groups = B >> 3;
        int my_inner = 0, my_prims = 0;
        for (int base = c0; base < c1; base += groups) {
            int t = base + group;
            bool valid = t < c1;
            int ni, np;
            int child = a.rec ? collapse_children_rec(a, cur[valid ? t : c0], j, ni, np) : collapse_children_g(a, cur[valid ? t : c0], j, ni, np);
            if (valid) {
                a.slots[8 * (size_t)t + j] = child;
                if (j == 0) {
                    a.counts[2 * (size_t)t] = ni;
                    a.counts[2 * (size_t)t + 1] = np;
                    my_inner += ni;
                    my_prims += np;
                }
            }
        }
        FINE(1)
        for (int o = 16; o; o >>= 1) {
            my_inner += __shfl_xor_sync(0xffffffffu, my_inner, o);
            my_prims += __shfl_xor_sync(0xffffffffu, my_prims, o);
        }
        if (lane == 0) { s_warp[0][warp] = my_inner; s_warp[1][warp] = my_prims; }
        __syncthreads();
        FINE(2)
        if (threadIdx.x == 0) {
            int ti = 0, tp = 0;
            for (int w = 0; w < B / 32; w++) { ti += s_warp[0][w]; tp += s_warp[1][w]; }
            btot[2 * bid] = ti;
            btot[2 * bid + 1] = tp;
        }
        sync();
        FINE(3)
        if (a.lvl_ns && bid == 0 && threadIdx.x == 0 && levels < 8) a.lvl_ns[2 * levels] = collapse_ns();
        // phase B: block offsets (every block sums the totals of the blocks before it), ordered scan, emit
        int pre_i = 0, pre_p = 0, tot_i = 0, tot_p = 0;
        for (int b = threadIdx.x; b < G; b += B) {
            int vi = btot[2 * b], vp = btot[2 * b + 1];
            tot_i += vi; tot_p += vp;
            if (b < bid) { pre_i += vi; pre_p += vp; }
        }
        for (int o = 16; o; o >>= 1) {
            pre_i += __shfl_xor_sync(0xffffffffu, pre_i, o);
            pre_p += __shfl_xor_sync(0xffffffffu, pre_p, o);
            tot_i += __shfl_xor_sync(0xffffffffu, tot_i, o);
            tot_p += __shfl_xor_sync(0xffffffffu, tot_p, o);
        }
        __syncthreads();
        if (lane == 0) { s_warp[0][warp] = pre_i; s_warp[1][warp] = pre_p; }
        __syncthreads();
        if (threadIdx.x == 0) {
            int pi = 0, pp = 0;
            for (int w = 0; w < B / 32; w++) { pi += s_warp[0][w]; pp += s_warp[1][w]; }
            s_carry[0] = pi; s_carry[1] = pp;
        }
        __syncthreads();
        if (lane == 0) { s_warp[0][warp] = tot_i; s_warp[1][warp] = tot_p; }
        __syncthreads();
        if (threadIdx.x == 0) {
            int ti = 0, tp = 0;
            for (int w = 0; w < B / 32; w++) { ti += s_warp[0][w]; tp += s_warp[1][w]; }
            s_base[0] = ti; s_base[1] = tp;
        }
        __syncthreads();
        FINE(4)
        const int level_inner = s_base[0], level_prims = s_base[1];
        const int next_base = level_base + level_n;
        for (int t0 = c0; t0 < c1; t0 += B) {
            int t = t0 + threadIdx.x;
            int ni = 0, np = 0;
            if (t < c1) { ni = a.counts[2 * (size_t)t]; np = a.counts[2 * (size_t)t + 1]; }
            int xi = ni, xp = np;
            for (int o = 1; o < 32; o <<= 1) {
                int yi = __shfl_up_sync(0xffffffffu, xi, o), yp = __shfl_up_sync(0xffffffffu, xp, o);
                if (lane >= o) { xi += yi; xp += yp; }
            }
            __syncthreads();
            if (lane == 31) { s_warp[0][warp] = xi; s_warp[1][warp] = xp; }
            __syncthreads();
            int wi = 0, wp = 0;
            for (int w = 0; w < warp; w++) { wi += s_warp[0][w]; wp += s_warp[1][w]; }
            int ci = s_carry[0], cp = s_carry[1];
            if (t < c1) {
                // child_base indexes nodes8 (absolute); prim_base indexes prims8
                a.counts[2 * (size_t)t] = next_base + (ci + wi + xi - ni);
                a.counts[2 * (size_t)t + 1] = prim_total + cp + wp + xp - np;
            }
            __syncthreads();
            if (threadIdx.x == B - 1) { s_carry[0] = ci + wi + xi; s_carry[1] = cp + wp + xp; }
            __syncthreads();
        }
        __syncthreads();
        FINE(5)
        for (int base = c0; base < c1; base += groups) {
            int t = base + group;
            bool valid = t < c1;
            int tt = valid ? t : c0;
            int child_base = a.counts[2 * (size_t)tt];
            write_node8_g<true>(a, cur[tt], level_base + tt, a.slots[8 * (size_t)tt + j], j, child_base, a.counts[2 * (size_t)tt + 1],
                          next + (child_base - next_base), valid);
        }
        FINE(6)
        level_base = next_base;
        prim_total += level_prims;
        level_n = level_inner;
        levels++;
        if (level_n > max_width) max_width = level_n;
        sync();
#ifdef RT_FINE
        if (a.lvl_ns && bid == 0 && threadIdx.x == 0 && levels == 3) a.lvl_ns[16 + 7] = collapse_ns();
#endif
        if (a.lvl_ns && bid == 0 && threadIdx.x == 0 && levels <= 8) a.lvl_ns[2 * levels - 1] = collapse_ns();
    }
    if (bid == 0 && threadIdx.x == 0) {
        a.result[0] = level_base;
        a.result[1] = prim_total;
        a.result[2] = levels;
        a.result[3] = max_width;
    }
}

__global__ void __launch_bounds__(kThreads) k_collapse(CollapseArgs a) {
    GridSync sync;
    sync.bar = (unsigned*)(a.result + 8);  // zeroed by the host before the launch
    sync.n_blocks = gridDim.x;
    sync.epoch = 0;
    collapse_levels(a, sync, (int)gridDim.x, (int)blockIdx.x);
}

// Refit of the BVH8 (rt_bvh_refit): 8 lanes per node requantise the slots' boxes from the refitted BVH2
// boxes and re-emit the leaf triangles from the moved world-space triangles. Same code as the build's
// emit step, so a refit of an unmoved scene reproduces the built tree bit for bit.
__global__ void __launch_bounds__(kThreads) k_refit8(CollapseArgs a, int n_nodes8) {
    const int group = (blockIdx.x * blockDim.x + threadIdx.x) >> 3, j = threadIdx.x & 7;
    const bool valid = group < n_nodes8;
    const int node = valid ? group : 0;
    const unsigned* old = (const unsigned*)(a.nodes8 + 5 * (size_t)node);
    const int child_base = (int)old[4], prim_base = (int)old[5];
    write_node8_g<false>(a, a.node8_root[node], node, a.node8_slot[8 * (size_t)node + j], j, child_base, prim_base, nullptr, valid);
}


// ================================================================ one-launch build on one thread-block cluster
// Small scenes (n <= G * 4096 triangles; world16 has 17 424): every step of the pipeline above in ONE kernel on ONE
// cluster of G = 8 or 16 CTAs x 512 threads. The steps are separated by the hardware cluster barrier
// (barrier.cluster, ~0.2 us, release / acquire at cluster scope) instead of 14 kernel boundaries, three software grid
// barriers per collapse level through a global atomic and the chained global arrival counters of k_refit:
//   * scene bounds and the radix histograms are exchanged through distributed shared memory (map_shared_rank);
//   * the radix sort keeps k_radix_scatter's stable match_any ranking, one contiguous run of keys per warp;
//   * the BVH2 boxes of nodes whose leaves lie inside one CTA's chunk are united with arrival counters in SHARED
//     memory; the few nodes that span chunks take the union of a suffix of the first chunk, the whole chunks between and
//     a prefix of the last one (min / max are exact, so any association gives the same bits) — no dependency chain;
//   * the BVH8 collapse is collapse_levels() with the cluster barrier.
// Arrays stay in global memory (L2-resident, ~250 cycles) so that rt_bvh_download, rt_bvh_refit and the traversal see
// exactly what the multi-kernel path leaves; results are bit-identical to it and to oracle/lbvh_host.c.
constexpr int kCT = 512;                        // threads per CTA (128 registers each)
constexpr int kCW = kCT / 32;                   // warps per CTA
constexpr int kMaxIpw = 8;                      // 32-key runs per warp and radix pass
constexpr int kChunkCap = kCW * 32 * kMaxIpw;   // keys / leaves per CTA: 4096
constexpr int kClusterDynSmem = 160 * 1024;     // dynamic shared memory per CTA: 40 960 keys (larger scenes read global memory)

struct ClusterBuildArgs {
    SceneView s;
    InstRec* insts;            // instance boxes are written here (create_boxes)
    const float* mesh_box;
    int n_insts, do_inst_boxes;
    const int* prim_offset;
    int n;
    float4* prim_world;
    float* prim_box;
    int* bounds_i;
    float* bounds_f;
    unsigned long long* keys_a; unsigned long long* keys_b;
    int* vals_a; int* vals_b;
    int key_shift, n_passes;
    int* left; int* right; int* parent; int* range_lo; int* range_hi;
    float* node_box;
    float* edge_box;           // 6 floats x (n prefix unions, n suffix unions, G chunk totals)
    uint4* rec;                // 4 x uint4 per BVH2 node (collapse_children_rec), or null
    CollapseArgs c;
    unsigned long long* phase_ns;  // 12 time stamps of CTA 0 (RT_BUILD_TIMERS), else null
    int smem_words;                // dynamic shared memory, in 32-bit words
};

struct ClusterSync {
    __device__ __forceinline__ void operator()() { cg::this_cluster().sync(); }
};

__device__ __forceinline__ unsigned long long build_global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void box_unite(float* acc, const float* b) {
#pragma unroll
    for (int k = 0; k < 3; k++) { acc[k] = fminf(acc[k], b[k]); acc[3 + k] = fmaxf(acc[3 + k], b[3 + k]); }
}

__global__ void __launch_bounds__(kCT, 1) k_build_cluster(ClusterBuildArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const int crank = (int)cluster.block_rank(), G = (int)cluster.num_blocks();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int T = G * kCT, gt = crank * kCT + tid;
    const int n = a.n;
    __shared__ unsigned s_big[kChunkCap];  // radix: digit counts per warp [kCW][kBins]; refit: arrival counters of the chunk
    __shared__ unsigned s_hist[kBins];     // this CTA's digit totals, read by every CTA of the cluster
    __shared__ int s_bounds[6];            // this CTA's scene bounds, read by every CTA of the cluster
    __shared__ int s_red[6][kCW];
    __shared__ unsigned s_wtot[kCW];
    __shared__ float s_scan[2][kCW][6];
    __shared__ float s_gb[6];
    // dynamic shared memory (a.smem_words 32-bit words): first the per-instance triangle prefix for flatten's binary
    // search, later the sorted 32-bit keys for Karras' binary searches — dependent loads at 29 cycles instead of an L2
    // round trip each (flatten 11 of them per triangle, Karras ~2 log2 n per node)
    extern __shared__ unsigned s_dyn[];
    const bool clock = a.phase_ns != nullptr && crank == 0 && tid == 0;
    if (clock) a.phase_ns[0] = build_global_ns();
    const bool offs_in_smem = a.n_insts + 1 <= a.smem_words;
    if (offs_in_smem) {
        for (int i = tid; i <= a.n_insts; i += kCT) s_dyn[i] = (unsigned)a.prim_offset[i];
        __syncthreads();
    }
    const int* prim_offset = offs_in_smem ? (const int*)s_dyn : a.prim_offset;

    // ---- create_boxes, world-space triangles, scene bounds
    if (a.do_inst_boxes)
        for (int i = gt; i < a.n_insts; i += T) {
            InstRec I = a.insts[i];
            instance_box(I, a.s.meshes, a.mesh_box);
#pragma unroll
            for (int k = 0; k < 6; k++) a.insts[i].box[k] = I.box[k];
            a.insts[i].box_ok = I.box_ok;
        }
    {
        int bl[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, bh[3] = {(int)0x80000000, (int)0x80000000, (int)0x80000000};
        for (int p = gt; p < n; p += T) {
            float lo[3], hi[3];
            flatten_prim(a.s, prim_offset, p, a.prim_world, a.prim_box, lo, hi);
#pragma unroll
            for (int k = 0; k < 3; k++) { bl[k] = min(bl[k], float_flip(lo[k])); bh[k] = max(bh[k], float_flip(hi[k])); }
        }
#pragma unroll
        for (int k = 0; k < 3; k++) {
            int l = bl[k], h = bh[k];
            for (int o = 16; o; o >>= 1) {
                l = min(l, __shfl_xor_sync(0xffffffffu, l, o));
                h = max(h, __shfl_xor_sync(0xffffffffu, h, o));
            }
            if (lane == 0) { s_red[k][warp] = l; s_red[3 + k][warp] = h; }
        }
        __syncthreads();
        if (tid < 6) {
            int v = s_red[tid][0];
            for (int w = 1; w < kCW; w++) v = tid < 3 ? min(v, s_red[tid][w]) : max(v, s_red[tid][w]);
            s_bounds[tid] = v;
        }
    }
    cluster.sync();
    if (tid < 6) {
        int v = s_bounds[tid];
        for (int r = 0; r < G; r++) {
            const int* rb = cluster.map_shared_rank(s_bounds, r);
            v = tid < 3 ? min(v, rb[tid]) : max(v, rb[tid]);
        }
        s_gb[tid] = float_unflip(v);
        if (crank == 0) { a.bounds_i[tid] = v; a.bounds_f[tid] = float_unflip(v); }
    }
    __syncthreads();
    if (clock) a.phase_ns[1] = build_global_ns();

    // ---- Morton keys
    {
        const float blo[3] = {s_gb[0], s_gb[1], s_gb[2]}, bhi[3] = {s_gb[3], s_gb[4], s_gb[5]};
        for (int p = gt; p < n; p += T) {
            a.keys_a[p] = morton_key(a.prim_world, p, blo, bhi, a.key_shift);
            a.vals_a[p] = p;
        }
    }
    cluster.sync();
    if (clock) a.phase_ns[2] = build_global_ns();

    // ---- stable LSD radix sort. Warp gw owns the ipw consecutive 32-key runs [gw * ipw, (gw + 1) * ipw): element order =
    // (CTA, warp, run, lane), which is index order.
    unsigned long long* kin = a.keys_a;
    unsigned long long* kout = a.keys_b;
    int* vin = a.vals_a;
    int* vout = a.vals_b;
    {
        const int nruns = (n + 31) >> 5, nwarps = G * kCW;
        const int ipw = (nruns + nwarps - 1) / nwarps;  // <= kMaxIpw (host-checked)
        const int gw = crank * kCW + warp;
        for (int pass = 0; pass < a.n_passes; pass++) {
            const int shift = pass * 8;
            for (int i = tid; i < kCW * kBins; i += kCT) s_big[i] = 0u;
            __syncthreads();
            unsigned long long k[kMaxIpw];
            int v[kMaxIpw];
#pragma unroll
            for (int it = 0; it < kMaxIpw; it++) {
                k[it] = ~0ull; v[it] = 0;
                if (it < ipw) {
                    const int idx = ((gw * ipw + it) << 5) + lane;
                    const bool ok = idx < n;
                    if (ok) { k[it] = kin[idx]; v[it] = vin[idx]; }
                    const unsigned d = (unsigned)(k[it] >> shift) & 0xffu;
                    const unsigned act = __ballot_sync(0xffffffffu, ok);
                    if (ok) {
                        const unsigned peers = __match_any_sync(act, d);
                        if (lane == (31 - __clz((int)peers))) s_big[warp * kBins + d] += (unsigned)__popc(peers);
                    }
                    __syncwarp();
                }
            }
            __syncthreads();
            if (tid < kBins) {
                unsigned t = 0;
                for (int w = 0; w < kCW; w++) t += s_big[w * kBins + tid];
                s_hist[tid] = t;
            }
            cluster.sync();  // every CTA's digit totals are in its s_hist
            unsigned tot = 0, before = 0;
            if (tid < kBins)
                for (int r = 0; r < G; r++) {
                    const unsigned* rh = cluster.map_shared_rank(s_hist, r);
                    const unsigned c = rh[tid];
                    tot += c;
                    if (r < crank) before += c;
                }
            unsigned x = tot;  // exclusive scan over the digits (threads 0..255)
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += y;
            }
            if (lane == 31) s_wtot[warp] = x;
            __syncthreads();
            if (tid < kBins) {
                unsigned run = x - tot + before;
                for (int w = 0; w < warp; w++) run += s_wtot[w];
                for (int w = 0; w < kCW; w++) {  // counts -> running offsets, warp by warp
                    const unsigned c = s_big[w * kBins + tid];
                    s_big[w * kBins + tid] = run;
                    run += c;
                }
            }
            __syncthreads();
#pragma unroll
            for (int it = 0; it < kMaxIpw; it++) {
                if (it < ipw) {
                    const int idx = ((gw * ipw + it) << 5) + lane;
                    const bool ok = idx < n;
                    const unsigned d = (unsigned)(k[it] >> shift) & 0xffu;
                    const unsigned act = __ballot_sync(0xffffffffu, ok);
                    if (ok) {
                        const unsigned peers = __match_any_sync(act, d);
                        const unsigned rank = (unsigned)__popc(peers & ((1u << lane) - 1u));
                        const unsigned pos = s_big[warp * kBins + d] + rank;
                        kout[pos] = k[it];
                        vout[pos] = v[it];
                        __syncwarp(act);
                        if (lane == (31 - __clz((int)peers))) s_big[warp * kBins + d] += (unsigned)__popc(peers);
                    }
                    __syncwarp();
                }
            }
            cluster.sync();  // the pass is in global memory; s_hist may be rewritten
            unsigned long long* tk = kin; kin = kout; kout = tk;
            int* tv = vin; vin = vout; vout = tv;
        }
    }
    if (clock) a.phase_ns[3] = build_global_ns();

    // ---- Karras topology
    if (a.key_shift == 31 && n <= a.smem_words) {
        for (int i = tid; i < n; i += kCT) s_dyn[i] = (unsigned)kin[i];  // every CTA holds all keys
        __syncthreads();
        for (int i = gt; i < n - 1; i += T) karras_node((const unsigned*)s_dyn, n, i, a.left, a.right, a.parent, a.range_lo, a.range_hi);
    } else {
        for (int i = gt; i < n - 1; i += T) karras_node(kin, n, i, a.left, a.right, a.parent, a.range_lo, a.range_hi);
    }
    cluster.sync();
    if (clock) a.phase_ns[4] = build_global_ns();

    // ---- BVH2 boxes
    const int per = (n + G - 1) / G;  // leaves per CTA, <= kChunkCap
    const int L0 = min(n, crank * per), L1 = min(n, L0 + per);
    float* pre = a.edge_box;
    float* suf = a.edge_box + 6 * (size_t)n;
    float* ctot = a.edge_box + 12 * (size_t)n;
    for (int i = tid; i < per; i += kCT) s_big[i] = 0u;
    // The climb walks one tree level per step, and every step is a chain of dependent loads (parent -> range and children ->
    // sibling box): in global memory three L2 round trips per level (38 us for world16's 15 levels). When the chunk's topology
    // and boxes fit the CTA's dynamic shared memory (72 bytes per leaf) the whole climb runs there; global memory only gets the
    // finished boxes. A node whose leaves lie inside [L0, L1) has its own id and its children inside the chunk too (a Karras
    // node's id is one end of its range), so every index below stays local.
    const bool climb_smem = (size_t)per * 72 <= (size_t)a.smem_words * 4 && n > 1;
    int4* s_topo = (int4*)s_dyn;                        // left, right, lo, hi of internal node L0 + k
    int* s_pint = (int*)(s_topo + per);                 // parent of internal node L0 + k
    int* s_pleaf = s_pint + per;                        // parent of leaf L0 + k
    volatile float* s_bint = (volatile float*)(s_pleaf + per);  // box of internal node L0 + k
    volatile float* s_bleaf = s_bint + 6 * (size_t)per;         // box of leaf L0 + k
    if (climb_smem) {
        __syncthreads();  // Karras' keys in s_dyn are dead (cluster.sync above); reuse the memory
        for (int k = tid; k < L1 - L0; k += kCT) {
            const int i = L0 + k;
            if (i < n - 1) {
                s_topo[k] = make_int4(a.left[i], a.right[i], a.range_lo[i], a.range_hi[i]);
                s_pint[k] = a.parent[i];
            }
            s_pleaf[k] = a.parent[n - 1 + i];
        }
    }
    __syncthreads();
    if (climb_smem) {
        for (int j = L0 + tid; j < L1; j += kCT) {
            const int p = vin[j];
            const int k0 = j - L0;
            float box[6];
#pragma unroll
            for (int k = 0; k < 6; k++) box[k] = a.prim_box[6 * (size_t)p + k];
            int child = n - 1 + j;
#pragma unroll
            for (int k = 0; k < 6; k++) { a.node_box[6 * (size_t)child + k] = box[k]; s_bleaf[6 * k0 + k] = box[k]; }
            int cur = s_pleaf[k0];
            while (cur >= L0 && cur < L1) {
                const int kk = cur - L0;
                const int4 tp = s_topo[kk];
                if (tp.z < L0 || tp.w >= L1) break;  // spans chunks: united from prefix / suffix unions below
                __threadfence_block();               // this thread's box of `child` is visible to the CTA before it arrives
                if (atomicAdd(&s_big[kk], 1u) == 0u) break;  // first arrival leaves; the second unites
                __threadfence_block();
                const int sib = tp.x == child ? tp.y : tp.x;
                const volatile float* sb = sib >= n - 1 ? s_bleaf + 6 * (size_t)(sib - (n - 1) - L0) : s_bint + 6 * (size_t)(sib - L0);
                float o[6];
#pragma unroll
                for (int k = 0; k < 6; k++) o[k] = sb[k];
                box_unite(box, o);
#pragma unroll
                for (int k = 0; k < 6; k++) { s_bint[6 * kk + k] = box[k]; a.node_box[6 * (size_t)cur + k] = box[k]; }
                child = cur;
                cur = s_pint[kk];
            }
        }
    } else
    for (int j = L0 + tid; j < L1; j += kCT) {
        const int p = vin[j];
        float box[6];
#pragma unroll
        for (int k = 0; k < 6; k++) box[k] = a.prim_box[6 * (size_t)p + k];
        int child = n - 1 + j;
#pragma unroll
        for (int k = 0; k < 6; k++) a.node_box[6 * (size_t)child + k] = box[k];
        if (n == 1) continue;
        int cur = a.parent[child];
        while (cur >= 0) {
            const int lo = a.range_lo[cur], hi = a.range_hi[cur];
            if (lo < L0 || hi >= L1) break;  // spans chunks: united from prefix / suffix unions below
            __threadfence_block();           // this thread's box of `child` is visible to the CTA before it arrives
            if (atomicAdd(&s_big[cur - L0], 1u) == 0u) break;  // first arrival leaves; the second unites
            __threadfence_block();
            const int l = a.left[cur], r = a.right[cur];
            const volatile float* sb = a.node_box + 6 * (size_t)(l == child ? r : l);
            float o[6];
#pragma unroll
            for (int k = 0; k < 6; k++) o[k] = sb[k];
            box_unite(box, o);
#pragma unroll
            for (int k = 0; k < 6; k++) a.node_box[6 * (size_t)cur + k] = box[k];
            child = cur;
            cur = a.parent[cur];
        }
    }
    __syncthreads();
    {   // prefix and suffix unions of the chunk's leaf boxes: thread t owns the R consecutive leaves [j0, j1)
        const int R = (per + kCT - 1) / kCT;
        const int j0 = min(L1, L0 + tid * R), j1 = min(L1, j0 + R);
        float run[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
        for (int j = j0; j < j1; j++) box_unite(run, a.node_box + 6 * (size_t)(n - 1 + j));
        float inc[6], dec[6];
#pragma unroll
        for (int k = 0; k < 6; k++) { inc[k] = run[k]; dec[k] = run[k]; }
        for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
            for (int k = 0; k < 6; k++) {
                const float yu = __shfl_up_sync(0xffffffffu, inc[k], o), yd = __shfl_down_sync(0xffffffffu, dec[k], o);
                if (lane >= o) inc[k] = k < 3 ? fminf(inc[k], yu) : fmaxf(inc[k], yu);
                if (lane + o < 32) dec[k] = k < 3 ? fminf(dec[k], yd) : fmaxf(dec[k], yd);
            }
        }
        if (lane == 31) {
#pragma unroll
            for (int k = 0; k < 6; k++) s_scan[0][warp][k] = inc[k];
        }
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < 6; k++) s_scan[1][warp][k] = dec[k];
        }
        __syncthreads();
        float ex[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
        float sx[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
        for (int w = 0; w < warp; w++) box_unite(ex, s_scan[0][w]);
        for (int w = warp + 1; w < kCW; w++) box_unite(sx, s_scan[1][w]);
#pragma unroll
        for (int k = 0; k < 6; k++) {
            const float yu = __shfl_up_sync(0xffffffffu, inc[k], 1), yd = __shfl_down_sync(0xffffffffu, dec[k], 1);
            if (lane > 0) ex[k] = k < 3 ? fminf(ex[k], yu) : fmaxf(ex[k], yu);
            if (lane < 31) sx[k] = k < 3 ? fminf(sx[k], yd) : fmaxf(sx[k], yd);
        }
        for (int j = j0; j < j1; j++) {
            box_unite(ex, a.node_box + 6 * (size_t)(n - 1 + j));
#pragma unroll
            for (int k = 0; k < 6; k++) pre[6 * (size_t)j + k] = ex[k];
        }
        for (int j = j1 - 1; j >= j0; j--) {
            box_unite(sx, a.node_box + 6 * (size_t)(n - 1 + j));
#pragma unroll
            for (int k = 0; k < 6; k++) suf[6 * (size_t)j + k] = sx[k];
        }
        if (tid == 0) {
            float t6[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
            for (int w = 0; w < kCW; w++) box_unite(t6, s_scan[0][w]);
#pragma unroll
            for (int k = 0; k < 6; k++) ctot[6 * (size_t)crank + k] = t6[k];
        }
    }
    cluster.sync();
    if (clock) a.phase_ns[5] = build_global_ns();
    for (int i = gt; i < n - 1; i += T) {  // nodes whose leaves span chunks
        const int lo = a.range_lo[i], hi = a.range_hi[i];
        const int cl = lo / per, ch = hi / per;
        if (cl == ch) continue;
        float box[6];
#pragma unroll
        for (int k = 0; k < 6; k++) box[k] = suf[6 * (size_t)lo + k];
        box_unite(box, pre + 6 * (size_t)hi);
        for (int c = cl + 1; c < ch; c++) box_unite(box, ctot + 6 * (size_t)c);
#pragma unroll
        for (int k = 0; k < 6; k++) a.node_box[6 * (size_t)i + k] = box[k];
    }
    cluster.sync();
    if (a.rec) {  // per-node records for the collapse (collapse_children_rec)
        for (int id = gt; id < 2 * n - 1; id += T) {
            float box[6];
#pragma unroll
            for (int k = 0; k < 6; k++) box[k] = a.node_box[6 * (size_t)id + k];
            if (id >= n - 1) {
                rec_store(a.rec, id, box, -1, -1, id - (n - 1), id - (n - 1), -1.0f, -1.0f, 0, 0);
            } else {
                const int l = a.left[id], r = a.right[id];
                const int sl = l >= n - 1 ? 1 : a.range_hi[l] - a.range_lo[l] + 1;
                const int sr = r >= n - 1 ? 1 : a.range_hi[r] - a.range_lo[r] + 1;
                float bl[6], br[6];
#pragma unroll
                for (int k = 0; k < 6; k++) { bl[k] = a.node_box[6 * (size_t)l + k]; br[k] = a.node_box[6 * (size_t)r + k]; }
                rec_store(a.rec, id, box, l, r, a.range_lo[id], a.range_hi[id], sl <= 3 ? -1.0f : half_area(bl), sr <= 3 ? -1.0f : half_area(br), sl, sr);
            }
        }
        cluster.sync();
    }
    if (clock) a.phase_ns[6] = build_global_ns();

    // ---- BVH8
    ClusterSync csync;
    collapse_levels(a.c, csync, G, crank);
    if (clock) a.phase_ns[7] = build_global_ns();
}

}  // namespace

namespace rth {

static inline int grid_for(int n, int per = kThreads) { return n > 0 ? (n + per - 1) / per : 1; }

// Stable LSD radix sort of (key, value) pairs, 8 bits per pass over the low 8 * n_passes key bits, ping-ponging
// between the a and b buffers (swapped in place: after the call `keys_a` / `vals_a` point at the sorted data).
// hist: kBins * (tiles + 1) counters. Returns the number of kernels launched.
size_t radix_hist_bytes(int n) { return sizeof(unsigned) * kBins * ((size_t)((n + kRadixTile - 1) / kRadixTile) + 1); }
int radix_sort_pairs(cudaStream_t st, unsigned long long*& keys_a, unsigned long long*& keys_b, int*& vals_a, int*& vals_b,
                     unsigned* hist, int n, int n_passes) {
    const int nb = (n + kRadixTile - 1) / kRadixTile;
    const bool fused_scan = nb <= 64;
    int launches = 0;
    for (int pass = 0; pass < n_passes; pass++) {
        int shift = pass * 8;
        k_radix_hist<<<nb, kThreads, 0, st>>>(keys_a, n, shift, hist, nb);
        if (fused_scan) {
            k_radix_scatter<true><<<nb, kThreads, 0, st>>>(keys_a, vals_a, n, shift, hist, nb, keys_b, vals_b);
            launches += 2;
        } else {
            k_radix_scan<<<kBins, kThreads, 0, st>>>(hist, nb);
            k_radix_scatter<false><<<nb, kThreads, 0, st>>>(keys_a, vals_a, n, shift, hist, nb, keys_b, vals_b);
            launches += 3;
        }
        std::swap(keys_a, keys_b);
        std::swap(vals_a, vals_b);
    }
    return launches;
}

// Cluster size of the one-launch build for a scene of n triangles, or 0 to use the multi-kernel pipeline.
static int cluster_build_size(int n) {
    if (n < 1) return 0;
    int want = 16;
    if (const char* e = std::getenv("RT_BUILD_PATH")) { if (std::atoi(e) == 1) return 0; }
    if (const char* e = std::getenv("RT_BUILD_CLUSTER")) { int v = std::atoi(e); if (v == 8 || v == 16 || v == 4 || v == 2 || v == 1) want = v; }
    static int max_cluster = -1;  // largest cluster of k_build_cluster this device can co-schedule
    if (max_cluster < 0) {
        max_cluster = 0;
        cudaFuncSetAttribute(k_build_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        cudaFuncSetAttribute(k_build_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, kClusterDynSmem);
        for (int g = 16; g >= 1; g >>= 1) {
            cudaLaunchConfig_t cfg;
            std::memset(&cfg, 0, sizeof cfg);
            cfg.gridDim = dim3(g); cfg.blockDim = dim3(kCT); cfg.dynamicSmemBytes = kClusterDynSmem;
            cudaLaunchAttribute at;
            at.id = cudaLaunchAttributeClusterDimension;
            at.val.clusterDim.x = g; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
            cfg.attrs = &at; cfg.numAttrs = 1;
            int nc = 0;
            if (cudaOccupancyMaxActiveClusters(&nc, k_build_cluster, &cfg) == cudaSuccess && nc >= 1) { max_cluster = g; break; }
            (void)cudaGetLastError();
        }
    }
    int G = want < max_cluster ? want : max_cluster;
    if (G < 1) return 0;
    if ((long long)n > (long long)G * kChunkCap) return 0;
    return G;
}

static int launch_cluster_build(RtScene* sc, const SceneView& v, int n, int key_shift, int n_passes, int G) {
    Bvh& b = sc->bvh;
    cudaStream_t st = sc->stream;
    const size_t level_cap = (size_t)n + 8;
    static const bool use_rec = [] { const char* e = std::getenv("RT_BUILD_RECORDS"); return !(e && !std::atoi(e)); }();
    if (b.edge_box.reserve(sizeof(float) * 6 * (2 * (size_t)n + 16)) || (use_rec && b.rec.reserve(64 * 2 * (size_t)n)) ||
        b.collapse_scratch.reserve(sizeof(int) * (level_cap * 2 + level_cap * 8 + level_cap * 2 + (size_t)G * 4 + 16)))
        return RT_ERR_OUT_OF_MEMORY;
    int* bounds_i = b.scene_bounds.as<int>();
    ClusterBuildArgs a;
    std::memset(&a, 0, sizeof a);
    a.s = v;
    a.insts = sc->d_insts.as<InstRec>();
    a.mesh_box = sc->d_mesh_box.as<float>();
    a.n_insts = sc->n_instances;
    a.do_inst_boxes = sc->boxes_valid ? 0 : 1;
    a.prim_offset = b.work.as<int>();
    a.n = n;
    a.prim_world = b.prim_world.as<float4>();
    a.prim_box = b.prim_box.as<float>();
    a.bounds_i = bounds_i;
    a.bounds_f = (float*)(bounds_i + 8);
    a.keys_a = b.keys_a.as<unsigned long long>(); a.keys_b = b.keys_b.as<unsigned long long>();
    a.vals_a = b.vals_a.as<int>(); a.vals_b = b.vals_b.as<int>();
    a.key_shift = key_shift; a.n_passes = n_passes;
    a.left = b.left.as<int>(); a.right = b.right.as<int>(); a.parent = b.parent.as<int>();
    a.range_lo = b.range_lo.as<int>(); a.range_hi = b.range_hi.as<int>();
    a.node_box = b.node_box.as<float>();
    a.edge_box = b.edge_box.as<float>();
    a.rec = use_rec ? b.rec.as<uint4>() : nullptr;
    static const bool timers = [] { const char* e = std::getenv("RT_BUILD_TIMERS"); return e && std::atoi(e); }();
    a.phase_ns = timers ? (unsigned long long*)(bounds_i + 16) : nullptr;
    int* w = b.collapse_scratch.as<int>();
    CollapseArgs& c = a.c;
    c.n = n;
    c.left = a.left; c.right = a.right; c.range_lo = a.range_lo; c.range_hi = a.range_hi;
    c.node_box = a.node_box;
    c.sorted_prim = a.vals_a;  // even number of passes
    c.prim_world = a.prim_world;
    c.nodes8 = b.nodes8.as<uint4>();
    c.prims8 = b.prims8.as<float4>();
    c.prim_order8 = b.prim_order8.as<int>();
    c.level_a = w; w += level_cap;
    c.level_b = w; w += level_cap;
    c.slots = w; w += level_cap * 8;
    c.counts = w; w += level_cap * 2;
    c.block_tot = w; w += (size_t)G * 4;
    c.result = w;
    c.node8_root = b.node8_root.as<int>();
    c.node8_slot = b.node8_slot.as<int>();
    c.lvl_ns = timers ? (unsigned long long*)((char*)b.scene_bounds.p + 256) : nullptr;
    c.rec = a.rec;
    cudaLaunchConfig_t cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(G); cfg.blockDim = dim3(kCT); cfg.stream = st; cfg.dynamicSmemBytes = kClusterDynSmem;
    a.smem_words = kClusterDynSmem / 4;
    cudaLaunchAttribute at;
    at.id = cudaLaunchAttributeClusterDimension;
    at.val.clusterDim.x = G; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
    cfg.attrs = &at; cfg.numAttrs = 1;
    RT_CUDA(cudaLaunchKernelEx(&cfg, k_build_cluster, a));
    sc->kernel_launches++;
    sc->boxes_valid = true;
    b.cluster_size = G;
    RT_CUDA(cudaMemcpyAsync(sc->h_pinned + 32, c.result, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (a.phase_ns) RT_CUDA(cudaMemcpyAsync(sc->h_pinned + 50, a.phase_ns, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    return RT_OK;
}

// Common tail of both builders: the collapse result is on its way to h_pinned[32..36).
static int finish_enqueue(RtScene* sc, RtBuildStats* stats, bool same_layout, int launches0, bool sync) {
    Bvh& b = sc->bvh;
    cudaStream_t st = sc->stream;
    RT_CUDA(rth::record(sc, sc->ev[1], st));
    RT_CUDA(cudaGetLastError());
    b.valid = true;  // provisional until finish_build() has looked at the result
    b.topology_valid = true;
    if (!same_layout) {
        b.topo_meshes.resize((size_t)sc->n_instances);
        for (int i = 0; i < sc->n_instances; i++) b.topo_meshes[i] = sc->instances[i].mesh;
    }
    b.build_launches = sc->kernel_launches - launches0;
    if (!sync) return RT_OK;
    RT_CUDA(cudaStreamSynchronize(st));
    return finish_build(sc, stats);
}

int build_bvh(RtScene* sc, RtBuildStats* stats, bool sync) {
    Bvh& b = sc->bvh;
    // prefix of triangle counts per instance (host). It only depends on which mesh every instance uses: when that is
    // what the previous build saw, the prefix already on the device (b.work) is reused.
    bool same_layout = b.topology_valid && (int)b.topo_meshes.size() == sc->n_instances && b.work.p != nullptr;
    b.valid = false;
    b.topology_valid = false;
    cudaStream_t st = sc->stream;
    int launches0 = sc->kernel_launches;
    for (int i = 0; same_layout && i < sc->n_instances; i++) same_layout = b.topo_meshes[i] == sc->instances[i].mesh;
    std::vector<int> off;
    int n = b.n_prims;
    if (!same_layout) {
        off.resize((size_t)sc->n_instances + 1);
        long long total = 0;
        for (int i = 0; i < sc->n_instances; i++) {
            off[i] = (int)total;
            total += sc->mesh_tri_count[sc->instances[i].mesh];
            if (total > 0x3fffffffLL) { set_error("too many world-space triangles"); return RT_ERR_LIMIT; }
        }
        off[sc->n_instances] = (int)total;
        n = (int)total;
    }
    b.n_prims = n;
    b.n_bvh2_nodes = n > 0 ? 2 * n - 1 : 0;
    b.n_nodes8 = 0;
    if (stats) { stats->n_prims = n; stats->n_bvh2_nodes = b.n_bvh2_nodes; stats->n_bvh8_nodes = 0; stats->kernel_launches = 0; stats->ms_build = 0.f; }
    RT_CUDA(rth::record(sc, sc->ev[0], st));
    const int cluster_G = cluster_build_size(n);  // > 0: the one-launch build, which also computes the instance boxes
    if (!cluster_G) { if (int rc = update_instance_boxes(sc)) return rc; }
    if (n == 0) {
        b.valid = true;
        RT_CUDA(rth::record(sc, sc->ev[1], st));
        RT_CUDA(cudaStreamSynchronize(st));
        return RT_OK;
    }
    int nb = (n + kRadixTile - 1) / kRadixTile;
    int max_nodes8 = n;  // every BVH8 node owns >= 1 BVH2 internal node with > 3 triangles (or is the root)
    if (b.work.reserve(sizeof(int) * ((size_t)sc->n_instances + 1)) || b.prim_world.reserve(sizeof(float4) * 3 * (size_t)n) ||
        b.prim_box.reserve(sizeof(float) * 6 * (size_t)n) || b.scene_bounds.reserve(512) ||
        b.keys_a.reserve(8 * (size_t)n) || b.keys_b.reserve(8 * (size_t)n) || b.vals_a.reserve(4 * (size_t)n) ||
        b.vals_b.reserve(4 * (size_t)n) || b.hist.reserve(sizeof(unsigned) * kBins * ((size_t)nb + 1)) ||
        b.left.reserve(4 * (size_t)n) || b.right.reserve(4 * (size_t)n) || b.parent.reserve(4 * 2 * (size_t)n) ||
        b.range_lo.reserve(4 * (size_t)n) || b.range_hi.reserve(4 * (size_t)n) || b.flags.reserve(4 * (size_t)n) ||
        b.node_box.reserve(sizeof(float) * 6 * 2 * (size_t)n) || b.nodes8.reserve(sizeof(uint4) * 5 * (size_t)max_nodes8) ||
        b.prims8.reserve(sizeof(float4) * 3 * (size_t)n) || b.prim_order8.reserve(4 * (size_t)n) ||
        b.node8_root.reserve(4 * (size_t)max_nodes8) || b.node8_slot.reserve(4 * 8 * (size_t)max_nodes8))
        return RT_ERR_OUT_OF_MEMORY;
    if (!same_layout) {
        // staged through pinned memory (a pageable source makes the copy synchronous with the host: 1.2 MB at 314 k instances)
        const size_t bytes = sizeof(int) * off.size();
        if (b.h_off_cap < bytes) {
            if (b.h_off) { RT_CUDA(cudaStreamSynchronize(st)); RT_CUDA(cudaFreeHost(b.h_off)); b.h_off = nullptr; b.h_off_cap = 0; }
            RT_CUDA(cudaHostAlloc((void**)&b.h_off, bytes + bytes / 2, cudaHostAllocDefault));
            b.h_off_cap = bytes + bytes / 2;
        } else if (b.off_in_flight) {
            RT_CUDA(cudaEventSynchronize(sc->ev[11]));
        }
        std::memcpy(b.h_off, off.data(), bytes);
        RT_CUDA(cudaMemcpyAsync(b.work.p, b.h_off, bytes, cudaMemcpyHostToDevice, st));
        RT_CUDA(rth::record(sc, sc->ev[11], st));
        b.off_in_flight = true;
        b.prefix_uploads++;
    }
    SceneView v = make_view(sc, false);
    int* bounds_i = b.scene_bounds.as<int>();
    float* bounds_f = (float*)(bounds_i + 8);
    const int key_shift = n <= (1 << 20) ? 31 : 0;
    const int n_passes = key_shift ? 4 : 8;
    // One launch on one thread-block cluster when the scene fits (n <= G * 4096); RT_BUILD_PATH=1 forces the multi-kernel
    // pipeline, RT_BUILD_CLUSTER=8|16 the cluster size (read per build so that tests can switch).
    if (cluster_G) {
        if (int rc = launch_cluster_build(sc, v, n, key_shift, n_passes, cluster_G)) return rc;
        b.sorted_keys = b.keys_a.as<unsigned long long>();  // even number of passes: back in keys_a / vals_a
        b.sorted_vals = b.vals_a.as<int>();
        b.build_path = 2;
        return finish_enqueue(sc, stats, same_layout, launches0, sync);
    }
    k_bounds_init<<<1, 32, 0, st>>>(bounds_i);
    k_flatten<<<grid_for(n), kThreads, 0, st>>>(v, b.work.as<int>(), n, b.prim_world.as<float4>(), b.prim_box.as<float>(), bounds_i);
    k_morton<<<grid_for(n), kThreads, 0, st>>>(b.prim_box.as<float>(), b.prim_world.as<float4>(), n, bounds_i,
                                               b.keys_a.as<unsigned long long>(), b.vals_a.as<int>(), bounds_f, key_shift);
    sc->kernel_launches += 3;
    unsigned long long* kin = b.keys_a.as<unsigned long long>();
    unsigned long long* kout = b.keys_b.as<unsigned long long>();
    int* vin = b.vals_a.as<int>();
    int* vout = b.vals_b.as<int>();
    sc->kernel_launches += radix_sort_pairs(st, kin, kout, vin, vout, b.hist.as<unsigned>(), n, n_passes);
    b.sorted_keys = kin;  // even number of passes: back in keys_a / vals_a
    b.sorted_vals = vin;
    if (n > 1) {
        k_karras<<<grid_for(n - 1), kThreads, 0, st>>>(kin, n, b.left.as<int>(), b.right.as<int>(), b.parent.as<int>(),
                                                       b.range_lo.as<int>(), b.range_hi.as<int>());
        sc->kernel_launches++;
    }
    RT_CUDA(cudaMemsetAsync(b.flags.p, 0, 4 * (size_t)n, st));
    k_refit<<<grid_for(n), kThreads, 0, st>>>(vin, b.prim_box.as<float>(), n, b.left.as<int>(), b.right.as<int>(),
                                              b.parent.as<int>(), b.flags.as<unsigned>(), b.node_box.as<float>());
    sc->kernel_launches++;
    // collapse
    int blocks_per_sm = 0;
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_collapse, kThreads, 0));
    if (blocks_per_sm < 1) { set_error("k_collapse cannot be made resident"); return RT_ERR_CUDA; }
    int G = sc->sm_count * (blocks_per_sm > 2 ? 2 : blocks_per_sm);
    int want = (n / 4 * 8 + kThreads - 1) / kThreads;  // 8 lanes per BVH8 node, about n/4 nodes on the widest level
    if (want < 1) want = 1;
    if (G > want) G = want;
    {   // experiment knob: grid barriers get cheaper with fewer blocks, the per-level work slower
        static const int env_blocks = [] { const char* e = std::getenv("RT_COLLAPSE_BLOCKS"); return e ? std::atoi(e) : 0; }();
        if (env_blocks > 0 && env_blocks < G) G = env_blocks;
    }
    size_t level_cap = (size_t)n + 8;
    size_t wbytes = sizeof(int) * (level_cap * 2 + level_cap * 8 + level_cap * 2 + (size_t)G * 4 + 16);
    static_assert(sizeof(int) == 4, "int");
    if (b.collapse_scratch.reserve(wbytes)) return RT_ERR_OUT_OF_MEMORY;
    int* w = b.collapse_scratch.as<int>();
    CollapseArgs a;
    a.n = n;
    a.left = b.left.as<int>(); a.right = b.right.as<int>(); a.range_lo = b.range_lo.as<int>(); a.range_hi = b.range_hi.as<int>();
    a.node_box = b.node_box.as<float>();
    a.sorted_prim = vin;
    a.prim_world = b.prim_world.as<float4>();
    a.nodes8 = b.nodes8.as<uint4>();
    a.prims8 = b.prims8.as<float4>();
    a.prim_order8 = b.prim_order8.as<int>();
    a.level_a = w; w += level_cap;
    a.level_b = w; w += level_cap;
    a.slots = w; w += level_cap * 8;
    a.counts = w; w += level_cap * 2;
    a.block_tot = w; w += (size_t)G * 4;
    a.result = w;
    a.node8_root = b.node8_root.as<int>();
    a.node8_slot = b.node8_slot.as<int>();
    a.lvl_ns = nullptr;
    a.rec = nullptr;
    RT_CUDA(cudaMemsetAsync(a.result, 0, 16 * sizeof(int), st));
    void* args[] = {&a};
    RT_CUDA(cudaLaunchCooperativeKernel((void*)k_collapse, dim3(G), dim3(kThreads), args, 0, st));
    sc->kernel_launches++;
    RT_CUDA(cudaMemcpyAsync(sc->h_pinned + 32, a.result, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    b.build_path = 1;
    return finish_enqueue(sc, stats, same_layout, launches0, sync);
}

bool refit_possible(const RtScene* sc) {
    const Bvh& b = sc->bvh;
    if (!b.topology_valid || b.n_prims <= 0 || b.n_nodes8 <= 0 || (int)b.topo_meshes.size() != sc->n_instances) return false;
    for (int i = 0; i < sc->n_instances; i++)
        if (b.topo_meshes[i] != sc->instances[i].mesh) return false;
    return true;
}

// Replaces nothing in the reference (it rebuilds from scratch every frame, src/raytracer.cu:102-107); the
// per-frame alternative for scenes whose instances only move. The BVH2 topology, the
// BVH8 child / slot / triangle layout and the sort order stay; k_flatten recomputes the world-space
// triangles and their inflated boxes, k_refit the BVH2 boxes, k_refit8 the quantised BVH8 boxes and the
// leaf-ordered triangles. Traversal stays a conservative filter in front of the exact test, so frames are
// identical to those of a rebuilt tree; only the number of nodes visited can grow as the scene drifts.
int refit_bvh(RtScene* sc, RtBuildStats* stats, bool sync) {
    Bvh& b = sc->bvh;
    if (!refit_possible(sc)) { set_error("no BVH topology to refit (build first; instances must keep their meshes)"); return RT_ERR_NO_BVH; }
    cudaStream_t st = sc->stream;
    int launches0 = sc->kernel_launches;
    const int n = b.n_prims;
    b.valid = false;
    RT_CUDA(rth::record(sc, sc->ev[0], st));
    if (int rc = update_instance_boxes(sc)) return rc;
    SceneView v = make_view(sc, false);
    int* bounds_i = b.scene_bounds.as<int>();
    k_bounds_init<<<1, 32, 0, st>>>(bounds_i);
    k_flatten<<<grid_for(n), kThreads, 0, st>>>(v, b.work.as<int>(), n, b.prim_world.as<float4>(), b.prim_box.as<float>(), bounds_i);
    RT_CUDA(cudaMemsetAsync(b.flags.p, 0, 4 * (size_t)n, st));
    k_refit<<<grid_for(n), kThreads, 0, st>>>(b.sorted_vals, b.prim_box.as<float>(), n, b.left.as<int>(), b.right.as<int>(),
                                              b.parent.as<int>(), b.flags.as<unsigned>(), b.node_box.as<float>());
    CollapseArgs a;
    memset(&a, 0, sizeof a);
    a.n = n;
    a.left = b.left.as<int>(); a.right = b.right.as<int>(); a.range_lo = b.range_lo.as<int>(); a.range_hi = b.range_hi.as<int>();
    a.node_box = b.node_box.as<float>();
    a.sorted_prim = b.sorted_vals;
    a.prim_world = b.prim_world.as<float4>();
    a.nodes8 = b.nodes8.as<uint4>();
    a.prims8 = b.prims8.as<float4>();
    a.prim_order8 = b.prim_order8.as<int>();
    a.node8_root = b.node8_root.as<int>();
    a.node8_slot = b.node8_slot.as<int>();
    k_refit8<<<grid_for(b.n_nodes8 * 8), kThreads, 0, st>>>(a, b.n_nodes8);
    sc->kernel_launches += 4;
    RT_CUDA(rth::record(sc, sc->ev[1], st));
    RT_CUDA(cudaGetLastError());
    b.valid = true;
    b.build_launches = sc->kernel_launches - launches0;
    // finish_build() re-reads the collapse result of the build this topology came from (still in h_pinned + 32)
    if (!sync) return RT_OK;
    RT_CUDA(cudaStreamSynchronize(st));
    return finish_build(sc, stats);
}

int finish_build(RtScene* sc, RtBuildStats* stats) {
    Bvh& b = sc->bvh;
    if (b.n_prims == 0) return RT_OK;
    const int* result = (const int*)(sc->h_pinned + 32);
    b.n_nodes8 = result[0];
    if (result[1] != b.n_prims) { b.valid = false; set_error("BVH8 collapse emitted a wrong number of triangles"); return RT_ERR_CUDA; }
    // Traversal (rt_bvh.cuh bvh_query) pushes at most ONE stack entry per node it descends through — the remaining hit
    // children of that node's group — so the stack never holds more entries than the tree has levels. A tree deeper than
    // the stack is refused here instead of dropping groups (= missing hits) at run time. An LBVH over duplicate Morton
    // codes splits on the triangle index, so its depth is bounded by key bits + log2(duplicates), not by the scene size.
    // RT_TEST_STACK_LIMIT lowers the limit so that tests can reach this path with a real tree.
    int stack_limit = RT_STACK_SIZE;
    if (const char* e = std::getenv("RT_TEST_STACK_LIMIT")) { int v = std::atoi(e); if (v > 0 && v < stack_limit) stack_limit = v; }
    b.n_levels8 = result[2];
    if (result[2] > stack_limit) {
        b.valid = false;
        b.topology_valid = false;
        set_error("BVH8 has " + std::to_string(result[2]) + " levels, more than the traversal stack holds (" + std::to_string(stack_limit) + ")");
        return RT_ERR_LIMIT;
    }
    if (stats) {
        stats->n_levels8 = result[2];
        stats->n_bvh8_nodes = b.n_nodes8;
        stats->kernel_launches = b.build_launches;
        float ms = 0.f;
        RT_CUDA(cudaEventElapsedTime(&ms, sc->ev[0], sc->ev[1]));
        stats->ms_build = ms;
    }
    return RT_OK;
}

}  // namespace rth

extern "C" {

int rt_bvh_download(RtScene* sc, uint64_t* morton, int32_t* perm, int32_t* left, int32_t* right, int32_t* parent,
                    float* boxes, float* scene_bounds6) {
    if (!sc) { rth::set_error("null scene"); return RT_ERR_INVALID_ARG; }
    rth::Bvh& b = sc->bvh;
    if (!b.valid) { rth::set_error("no BVH built"); return RT_ERR_NO_BVH; }
    size_t n = (size_t)b.n_prims;
    if (n == 0) return RT_OK;
    cudaStream_t st = sc->stream;
    if (morton) RT_CUDA(cudaMemcpyAsync(morton, b.sorted_keys, 8 * n, cudaMemcpyDeviceToHost, st));
    if (perm) RT_CUDA(cudaMemcpyAsync(perm, b.sorted_vals, 4 * n, cudaMemcpyDeviceToHost, st));
    if (left && n > 1) RT_CUDA(cudaMemcpyAsync(left, b.left.p, 4 * (n - 1), cudaMemcpyDeviceToHost, st));
    if (right && n > 1) RT_CUDA(cudaMemcpyAsync(right, b.right.p, 4 * (n - 1), cudaMemcpyDeviceToHost, st));
    if (parent && n > 1) RT_CUDA(cudaMemcpyAsync(parent, b.parent.p, 4 * (2 * n - 1), cudaMemcpyDeviceToHost, st));
    if (boxes) RT_CUDA(cudaMemcpyAsync(boxes, b.node_box.p, 24 * (2 * n - 1), cudaMemcpyDeviceToHost, st));
    if (scene_bounds6) RT_CUDA(cudaMemcpyAsync(scene_bounds6, (const char*)b.scene_bounds.p + 32, 24, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}

int rt_bvh8_download(RtScene* sc, void* nodes80, int32_t max_nodes, int32_t* n_nodes, void* prims48, int32_t* prim_order,
                     int32_t max_prims, int32_t* n_prims) {
    if (!sc) { rth::set_error("null scene"); return RT_ERR_INVALID_ARG; }
    rth::Bvh& b = sc->bvh;
    if (!b.valid) { rth::set_error("no BVH built"); return RT_ERR_NO_BVH; }
    if (n_nodes) *n_nodes = b.n_nodes8;
    if (n_prims) *n_prims = b.n_prims;
    cudaStream_t st = sc->stream;
    if (nodes80) {
        if (max_nodes < b.n_nodes8) { rth::set_error("node buffer too small"); return RT_ERR_INVALID_ARG; }
        if (b.n_nodes8) RT_CUDA(cudaMemcpyAsync(nodes80, b.nodes8.p, 80 * (size_t)b.n_nodes8, cudaMemcpyDeviceToHost, st));
    }
    if (prims48 || prim_order) {
        if (max_prims < b.n_prims) { rth::set_error("triangle buffer too small"); return RT_ERR_INVALID_ARG; }
        if (prims48 && b.n_prims) RT_CUDA(cudaMemcpyAsync(prims48, b.prims8.p, 48 * (size_t)b.n_prims, cudaMemcpyDeviceToHost, st));
        if (prim_order && b.n_prims) RT_CUDA(cudaMemcpyAsync(prim_order, b.prim_order8.p, 4 * (size_t)b.n_prims, cudaMemcpyDeviceToHost, st));
    }
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}

}  // extern "C"
