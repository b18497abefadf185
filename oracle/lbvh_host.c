/* lbvh_host.c — CPU ORACLE (test infrastructure only): sequential host build of the LBVH and of
 * the compressed BVH8 the CUDA engine traverses.
 *
 * The reference has no counterpart of this structure (its BVH is an implicit binary tree over
 * instance boxes, src/rayopt/bvh.cu:43-91); BASELINE.json's north_star asks for a GPU LBVH build
 * that is "bit-exact against a host reference build of the same algorithm". This file is that
 * host build, written independently of the kernels in gpu-ray-tracer_b200/csrc/rt_bvh_build.cu:
 *   - Morton codes by a bit loop (the kernels use magic-number bit spreading),
 *   - ordering by a comparison sort on (code, index) (the kernels use an LSD radix sort),
 *   - topology by top-down recursive splitting on the highest differing bit of the augmented
 *     keys (the kernels use Karras' 2012 per-node binary searches) with Karras' node numbering,
 *   - boxes by a post-order union (the kernels use atomics-ordered bottom-up refit),
 *   - BVH8 collapse, slot assignment and quantisation by a breadth-first queue (the kernels use a
 *     cooperative level-synchronous pass with scans).
 * tests/test_gpu_parity.py::test_bvh_build_bit_exact_vs_host compares every array of both builds for equality.
 *
 * Floating-point rules shared with the kernels (both sides compile without FMA contraction):
 * every expression below is written operation by operation in the order documented in
 * DESIGN.md ("BVH build, bit-exact rules").
 */
#include "lbvh_host.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define THR 1E-5f

typedef struct { float x, y, z; } V3;
typedef struct { float i, j, k, r; } Q4;

/* ---- the reference's rigid-transform arithmetic (same restatement as rt_oracle.c) ---- */
static V3 v3(float x, float y, float z) { V3 v = {x, y, z}; return v; }
static V3 v3add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static V3 v3scale(float c, V3 v) { return v3(v.x * c, v.y * c, v.z * c); }
static float v3dot(V3 a, V3 b) { float s = 0; s += a.x * b.x; s += a.y * b.y; s += a.z * b.z; return s; }
static float v3len(V3 v) { return sqrtf(v3dot(v, v)); }
static V3 v3normalized(V3 v) { float len = v3len(v); if (len > THR) return v3scale(1 / len, v); return v3(0, 0, 0); }
static float q4dot(Q4 a, Q4 b) { float s = 0; s += a.i * b.i; s += a.j * b.j; s += a.k * b.k; s += a.r * b.r; return s; }
static Q4 q4normalized(Q4 q) {
    float len = sqrtf(q4dot(q, q));
    Q4 z = {0, 0, 0, 0};
    if (len > THR) { float inv = 1 / len; Q4 r = {q.i * inv, q.j * inv, q.k * inv, q.r * inv}; return r; }
    return z;
}
static Q4 q4inverse(Q4 q) {
    float sq = q4dot(q, q);
    Q4 z = {0, 0, 0, 0};
    if (sq < THR) return z;
    float inv = 1 / sq;
    Q4 r = {q.i * -inv, q.j * -inv, q.k * -inv, q.r * inv};
    return r;
}
static Q4 q4mul(Q4 a, Q4 o) {
    Q4 r;
    r.i = a.i * o.r + a.r * o.i + a.j * o.k - a.k * o.j;
    r.j = a.j * o.r + a.r * o.j + a.k * o.i - a.i * o.k;
    r.k = a.k * o.r + a.r * o.k + a.i * o.j - a.j * o.i;
    r.r = a.r * o.r - a.i * o.i - a.j * o.j - a.k * o.k;
    return r;
}
static V3 q4rotate(Q4 q, V3 v) {
    float length = v3len(v);
    Q4 qv = {v.x, v.y, v.z, 0};
    Q4 prod = q4mul(q4mul(q4normalized(q), qv), q4inverse(q));
    return v3scale(length, v3normalized(v3(prod.i, prod.j, prod.k)));
}
/* ref: src/rayprimitives/entity.cu:23-29 point_from_local */
static V3 point_from_local(const RtPose* p, V3 v) {
    Q4 q = {p->rot[0], p->rot[1], p->rot[2], p->rot[3]};
    return v3add(q4rotate(q4inverse(q), v), v3(p->pos[0], p->pos[1], p->pos[2]));
}

struct OrBvh {
    int n, n_nodes8, levels;
    float* prim_world;      /* 12 floats per prim, input order: v0.xyz, inst | e1.xyz, tri | e2.xyz, L */
    float* prim_box;        /* 6 per prim (inflated) */
    float bounds[6];
    uint64_t* morton;       /* sorted */
    int32_t* perm;          /* sorted position -> prim */
    int32_t *left, *right, *parent, *range_lo, *range_hi;
    float* node_box;        /* 6 per BVH2 node (2n-1) */
    uint32_t* nodes8;       /* 20 words per node */
    float* prims8;          /* 12 floats per prim, leaf order */
    int32_t* prim_order8;
};

static float fbits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static uint32_t ubits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

/* ---- flatten ---- */
static void flatten(OrBvh* b, const RtSceneDesc* d) {
    int p = 0;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int inst = 0; inst < d->n_instances; inst++) {
        const RtInstance* I = &d->instances[inst];
        const RtMesh* M = &d->meshes[I->mesh];
        if (M->kind == RT_HITABLE_SPHERE) { /* world centre, zero edges, L = +inf (the triangle filter passes it) */
            V3 c = point_from_local(&I->pose, point_from_local(&M->pose, v3(0, 0, 0)));
            float* o = b->prim_world + 12 * (size_t)p;
            int32_t ii = inst, tt = M->tri_begin;
            float cc[3] = {c.x, c.y, c.z}, tl[3], th[3];
            o[0] = c.x; o[1] = c.y; o[2] = c.z; memcpy(&o[3], &ii, 4);
            o[4] = 0; o[5] = 0; o[6] = 0; memcpy(&o[7], &tt, 4);
            o[8] = 0; o[9] = 0; o[10] = 0; o[11] = INFINITY;
            for (int k = 0; k < 3; k++) {
                tl[k] = cc[k] - M->radius;
                th[k] = cc[k] + M->radius;
                if (tl[k] < lo[k]) lo[k] = tl[k];
                if (th[k] > hi[k]) hi[k] = th[k];
            }
            float amax = fmaxf(fmaxf(fmaxf(fabsf(tl[0]), fabsf(th[0])), fmaxf(fabsf(tl[1]), fabsf(th[1]))),
                               fmaxf(fabsf(tl[2]), fabsf(th[2])));
            float m = (2e-4f + 4e-5f * (6.0f * M->radius)) + 1e-6f * amax;
            for (int k = 0; k < 3; k++) {
                b->prim_box[6 * (size_t)p + k] = tl[k] - m;
                b->prim_box[6 * (size_t)p + 3 + k] = th[k] + m;
            }
            p++;
            continue;
        }
        for (int tri = M->tri_begin; tri < M->tri_end; tri++, p++) {
            V3 w[3];
            for (int k = 0; k < 3; k++) {
                const float* v = d->vertices + 3 * (size_t)d->triangles[tri].v[k];
                w[k] = point_from_local(&I->pose, point_from_local(&M->pose, v3(v[0], v[1], v[2])));
            }
            float e1[3] = {w[1].x - w[0].x, w[1].y - w[0].y, w[1].z - w[0].z};
            float e2[3] = {w[2].x - w[0].x, w[2].y - w[0].y, w[2].z - w[0].z};
            float L = ((fabsf(e1[0]) + fabsf(e1[1])) + fabsf(e1[2])) + ((fabsf(e2[0]) + fabsf(e2[1])) + fabsf(e2[2]));
            float* o = b->prim_world + 12 * (size_t)p;
            int32_t ii = inst, tt = tri;
            o[0] = w[0].x; o[1] = w[0].y; o[2] = w[0].z; memcpy(&o[3], &ii, 4);
            o[4] = e1[0]; o[5] = e1[1]; o[6] = e1[2]; memcpy(&o[7], &tt, 4);
            o[8] = e2[0]; o[9] = e2[1]; o[10] = e2[2]; o[11] = L;
            float wx[3][3] = {{w[0].x, w[1].x, w[2].x}, {w[0].y, w[1].y, w[2].y}, {w[0].z, w[1].z, w[2].z}};
            float tl[3], th[3], amax = 0.0f;
            for (int k = 0; k < 3; k++) {
                tl[k] = fminf(fminf(wx[k][0], wx[k][1]), wx[k][2]);
                th[k] = fmaxf(fmaxf(wx[k][0], wx[k][1]), wx[k][2]);
                if (tl[k] < lo[k]) lo[k] = tl[k];
                if (th[k] > hi[k]) hi[k] = th[k];
            }
            amax = fmaxf(fmaxf(fmaxf(fabsf(tl[0]), fabsf(th[0])), fmaxf(fabsf(tl[1]), fabsf(th[1]))),
                         fmaxf(fabsf(tl[2]), fabsf(th[2])));
            float m = (2e-4f + 4e-5f * L) + 1e-6f * amax;
            for (int k = 0; k < 3; k++) {
                b->prim_box[6 * (size_t)p + k] = tl[k] - m;
                b->prim_box[6 * (size_t)p + 3 + k] = th[k] + m;
            }
        }
    }
    for (int k = 0; k < 3; k++) { b->bounds[k] = lo[k]; b->bounds[3 + k] = hi[k]; }
}

/* ---- Morton ---- */
static uint64_t morton_of(const OrBvh* b, int p) {
    const float* o = b->prim_world + 12 * (size_t)p;
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
        float a = o[k], bb = o[k] + o[4 + k], c = o[k] + o[8 + k];
        float lo = fminf(fminf(a, bb), c), hi = fmaxf(fmaxf(a, bb), c);
        float cen = (lo + hi) * 0.5f;
        float ext = b->bounds[3 + k] - b->bounds[k];
        float n = ext > 0.0f ? (cen - b->bounds[k]) / ext : 0.0f;
        float sc = n * 2097152.0f;
        sc = fminf(fmaxf(sc, 0.0f), 2097151.0f);
        q[k] = (uint32_t)sc;
    }
    uint64_t code = 0;
    for (int bit = 20; bit >= 0; bit--) {
        code = (code << 1) | ((q[0] >> bit) & 1u);
        code = (code << 1) | ((q[1] >> bit) & 1u);
        code = (code << 1) | ((q[2] >> bit) & 1u);
    }
    return code;
}

typedef struct { uint64_t key; int32_t idx; } KeyIdx;
static int cmp_keyidx(const void* a, const void* b) {
    const KeyIdx* x = (const KeyIdx*)a; const KeyIdx* y = (const KeyIdx*)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx ? 1 : 0);
}

/* ---- topology: recursive split on the highest differing bit of (code, position) ---- */
/* common-prefix length of the augmented 96-bit keys of sorted positions i and j */
static int common_prefix(const OrBvh* b, int i, int j) {
    uint64_t a = b->morton[i], c = b->morton[j];
    if (a != c) return __builtin_clzll(a ^ c);
    return 64 + __builtin_clz((uint32_t)(i ^ j));
}
typedef struct { int node, first, last; } Range;

static void build_topology(OrBvh* b) {
    int n = b->n;
    if (n < 2) return;
    Range* stack = (Range*)malloc(sizeof(Range) * (size_t)(2 * n));
    int sp = 0;
    Range root = {0, 0, n - 1};
    stack[sp++] = root;
    b->parent[0] = -1;
    while (sp) {
        Range r = stack[--sp];
        int prefix = common_prefix(b, r.first, r.last);
        /* split = last position in [first, last) that shares more than `prefix` bits with `first` */
        int lo = r.first, hi = r.last;  /* invariant: cp(first, lo) > prefix, cp(first, hi) == prefix */
        while (hi - lo > 1) {
            int mid = lo + (hi - lo) / 2;
            if (common_prefix(b, r.first, mid) > prefix) lo = mid; else hi = mid;
        }
        int split = lo;
        int lc = (split == r.first) ? (n - 1 + split) : split;
        int rc = (split + 1 == r.last) ? (n - 1 + split + 1) : (split + 1);
        b->left[r.node] = lc; b->right[r.node] = rc;
        b->parent[lc] = r.node; b->parent[rc] = r.node;
        b->range_lo[r.node] = r.first; b->range_hi[r.node] = r.last;
        if (lc < n - 1) { Range c = {lc, r.first, split}; stack[sp++] = c; }
        if (rc < n - 1) { Range c = {rc, split + 1, r.last}; stack[sp++] = c; }
    }
    free(stack);
}

/* ---- boxes: post-order union ---- */
static void build_boxes(OrBvh* b) {
    int n = b->n;
    for (int j = 0; j < n; j++)
        memcpy(b->node_box + 6 * (size_t)(n - 1 + j), b->prim_box + 6 * (size_t)b->perm[j], 24);
    if (n < 2) return;
    /* internal nodes: children of node i always have larger range nesting; process by explicit post-order */
    int* order = (int*)malloc(sizeof(int) * (size_t)n);
    int* stack = (int*)malloc(sizeof(int) * (size_t)n);
    int cnt = 0, sp = 0;
    stack[sp++] = 0;
    while (sp) {
        int i = stack[--sp];
        order[cnt++] = i;
        if (b->left[i] < n - 1) stack[sp++] = b->left[i];
        if (b->right[i] < n - 1) stack[sp++] = b->right[i];
    }
    for (int k = cnt - 1; k >= 0; k--) {  /* reverse pre-order visits children before parents */
        int i = order[k];
        const float* l = b->node_box + 6 * (size_t)b->left[i];
        const float* r = b->node_box + 6 * (size_t)b->right[i];
        float* o = b->node_box + 6 * (size_t)i;
        for (int a = 0; a < 3; a++) { o[a] = fminf(l[a], r[a]); o[3 + a] = fmaxf(l[3 + a], r[3 + a]); }
    }
    free(order);
    free(stack);
}

/* ---- BVH8 collapse ---- */
static int node_size(const OrBvh* b, int id) { return id >= b->n - 1 ? 1 : b->range_hi[id] - b->range_lo[id] + 1; }
static int node_first(const OrBvh* b, int id) { return id >= b->n - 1 ? id - (b->n - 1) : b->range_lo[id]; }
static float half_area(const float* x) {
    float dx = x[3] - x[0], dy = x[4] - x[1], dz = x[5] - x[2];
    return (dx * dy + dy * dz) + dz * dx;
}

static void choose_children(const OrBvh* b, int root, int slot_child[8]) {
    int ch[8], nc = 0;
    if (root >= b->n - 1 || node_size(b, root) <= 3) {
        ch[nc++] = root;
    } else {
        ch[nc++] = b->left[root];
        ch[nc++] = b->right[root];
        while (nc < 8) {
            int best = -1;
            float best_area = -1.0f;
            for (int i = 0; i < nc; i++) {
                if (node_size(b, ch[i]) <= 3) continue;
                float ar = half_area(b->node_box + 6 * (size_t)ch[i]);
                if (ar > best_area) { best_area = ar; best = i; }
            }
            if (best < 0) break;
            int c = ch[best];
            ch[best] = b->left[c];
            ch[nc++] = b->right[c];
        }
    }
    const float* nb = b->node_box + 6 * (size_t)root;
    float ncx = (nb[0] + nb[3]) * 0.5f, ncy = (nb[1] + nb[4]) * 0.5f, ncz = (nb[2] + nb[5]) * 0.5f;
    float cost[8][8];
    for (int i = 0; i < nc; i++) {
        const float* cb = b->node_box + 6 * (size_t)ch[i];
        float dx = (cb[0] + cb[3]) * 0.5f - ncx, dy = (cb[1] + cb[4]) * 0.5f - ncy, dz = (cb[2] + cb[5]) * 0.5f - ncz;
        for (int s = 0; s < 8; s++) cost[i][s] = (((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy)) + ((s & 4) ? dz : -dz);
    }
    for (int s = 0; s < 8; s++) slot_child[s] = -1;
    int child_done[8] = {0}, slot_done[8] = {0};
    for (int it = 0; it < nc; it++) {
        int bi = -1, bs = -1;
        float bc = -INFINITY;
        for (int i = 0; i < nc; i++) {
            if (child_done[i]) continue;
            for (int s = 0; s < 8; s++) {
                if (slot_done[s]) continue;
                if (bi < 0 || cost[i][s] > bc) { bc = cost[i][s]; bi = i; bs = s; }
            }
        }
        child_done[bi] = 1; slot_done[bs] = 1;
        slot_child[bs] = ch[bi];
    }
}

static uint32_t pack4(const uint32_t* x) { return x[0] | (x[1] << 8) | (x[2] << 16) | (x[3] << 24); }

static void collapse(OrBvh* b) {
    int n = b->n;
    int* queue = (int*)malloc(sizeof(int) * (size_t)(n + 8));  /* BVH2 id of BVH8 node i, breadth first */
    int qh = 0, qt = 0, prim_total = 0, levels = 0;
    queue[qt++] = 0;
    while (qh < qt) {
        int level_end = qt;  /* nodes [qh, level_end) form one level; their inner children are appended in order */
        for (int idx = qh; idx < level_end; idx++) {
            int root = queue[idx];
            int sc[8];
            choose_children(b, root, sc);
            const float* nb = b->node_box + 6 * (size_t)root;
            float org[3] = {nb[0], nb[1], nb[2]}, scale[3], inv_scale[3];
            uint32_t eb[3];
            for (int k = 0; k < 3; k++) {
                float ext = nb[3 + k] - nb[k];
                float f = ext / 255.0f;
                uint32_t bits = ubits(f), e = (bits >> 23) & 0xffu;
                if (bits & 0x7fffffu) e += 1u;
                if (e < 1u) e = 1u;
                if (e > 253u) e = 253u;
                float s = fbits(e << 23);
                if (org[k] + 255.0f * s < nb[3 + k] && e < 253u) { e += 1u; s = fbits(e << 23); }
                eb[k] = e; scale[k] = s; inv_scale[k] = fbits((254u - e) << 23);
            }
            uint32_t imask = 0, meta[8], qlo[3][8], qhi[3][8];
            int child_base = qt, prim_base = prim_total, prim_off = 0;
            for (int s = 0; s < 8; s++) {
                meta[s] = 0;
                for (int k = 0; k < 3; k++) { qlo[k][s] = 0; qhi[k][s] = 0; }
                int c = sc[s];
                if (c < 0) continue;
                const float* cb = b->node_box + 6 * (size_t)c;
                for (int k = 0; k < 3; k++) {
                    float ql = floorf((cb[k] - org[k]) * inv_scale[k]);
                    ql = fminf(fmaxf(ql, 0.0f), 255.0f);
                    while (ql > 0.0f && org[k] + ql * scale[k] > cb[k]) ql -= 1.0f;
                    float qh2 = ceilf((cb[3 + k] - org[k]) * inv_scale[k]);
                    qh2 = fminf(fmaxf(qh2, 0.0f), 255.0f);
                    while (qh2 < 255.0f && org[k] + qh2 * scale[k] < cb[3 + k]) qh2 += 1.0f;
                    qlo[k][s] = (uint32_t)ql; qhi[k][s] = (uint32_t)qh2;
                }
                int sz = node_size(b, c);
                if (sz > 3) {
                    imask |= 1u << s;
                    meta[s] = 0x20u | (24u + (uint32_t)s);
                    queue[qt++] = c;
                } else {
                    uint32_t unary = sz == 1 ? 1u : (sz == 2 ? 3u : 7u);
                    meta[s] = (unary << 5) | (uint32_t)prim_off;
                    int first = node_first(b, c);
                    for (int j = 0; j < sz; j++) {
                        int dst = prim_base + prim_off + j;
                        b->prim_order8[dst] = first + j;
                        memcpy(b->prims8 + 12 * (size_t)dst, b->prim_world + 12 * (size_t)b->perm[first + j], 48);
                    }
                    prim_off += sz;
                }
            }
            prim_total += prim_off;
            uint32_t* o = b->nodes8 + 20 * (size_t)idx;
            o[0] = ubits(org[0]); o[1] = ubits(org[1]); o[2] = ubits(org[2]);
            o[3] = eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24);
            o[4] = (uint32_t)child_base; o[5] = (uint32_t)prim_base; o[6] = pack4(meta); o[7] = pack4(meta + 4);
            o[8] = pack4(qlo[0]); o[9] = pack4(qlo[0] + 4); o[10] = pack4(qlo[1]); o[11] = pack4(qlo[1] + 4);
            o[12] = pack4(qlo[2]); o[13] = pack4(qlo[2] + 4); o[14] = pack4(qhi[0]); o[15] = pack4(qhi[0] + 4);
            o[16] = pack4(qhi[1]); o[17] = pack4(qhi[1] + 4); o[18] = pack4(qhi[2]); o[19] = pack4(qhi[2] + 4);
        }
        qh = level_end;
        levels++;
    }
    b->n_nodes8 = qt;
    b->levels = levels;
    free(queue);
}

OrBvh* or_bvh_build(const RtSceneDesc* d) {
    OrBvh* b = (OrBvh*)calloc(1, sizeof(OrBvh));
    long long total = 0;
    for (int i = 0; i < d->n_instances; i++) {
        const RtMesh* M = &d->meshes[d->instances[i].mesh];
        total += M->tri_end - M->tri_begin;
    }
    int n = (int)total;
    b->n = n;
    size_t m = (size_t)(n > 0 ? n : 1);
    b->prim_world = (float*)calloc(12 * m, 4);
    b->prim_box = (float*)calloc(6 * m, 4);
    b->morton = (uint64_t*)calloc(m, 8);
    b->perm = (int32_t*)calloc(m, 4);
    b->left = (int32_t*)calloc(m, 4); b->right = (int32_t*)calloc(m, 4);
    b->parent = (int32_t*)calloc(2 * m, 4);
    b->range_lo = (int32_t*)calloc(m, 4); b->range_hi = (int32_t*)calloc(m, 4);
    b->node_box = (float*)calloc(12 * m, 4);
    b->nodes8 = (uint32_t*)calloc(20 * m, 4);
    b->prims8 = (float*)calloc(12 * m, 4);
    b->prim_order8 = (int32_t*)calloc(m, 4);
    if (n == 0) return b;
    flatten(b, d);
    KeyIdx* ki = (KeyIdx*)malloc(sizeof(KeyIdx) * (size_t)n);
    /* scenes up to 2^20 triangles keep the top 32 bits of the 63-bit code (four radix passes on the GPU) */
    int key_shift = n <= (1 << 20) ? 31 : 0;
    for (int p = 0; p < n; p++) { ki[p].key = morton_of(b, p) >> key_shift; ki[p].idx = p; }
    qsort(ki, (size_t)n, sizeof(KeyIdx), cmp_keyidx);
    for (int j = 0; j < n; j++) { b->morton[j] = ki[j].key; b->perm[j] = ki[j].idx; }
    free(ki);
    build_topology(b);
    build_boxes(b);
    collapse(b);
    return b;
}

void or_bvh_free(OrBvh* b) {
    if (!b) return;
    free(b->prim_world); free(b->prim_box); free(b->morton); free(b->perm); free(b->left); free(b->right);
    free(b->parent); free(b->range_lo); free(b->range_hi); free(b->node_box); free(b->nodes8); free(b->prims8);
    free(b->prim_order8); free(b);
}

int32_t or_bvh_n_prims(const OrBvh* b) { return b->n; }
int32_t or_bvh_n_nodes8(const OrBvh* b) { return b->n_nodes8; }
int32_t or_bvh_levels(const OrBvh* b) { return b->levels; }
const uint64_t* or_bvh_morton(const OrBvh* b) { return b->morton; }
const int32_t* or_bvh_perm(const OrBvh* b) { return b->perm; }
const int32_t* or_bvh_left(const OrBvh* b) { return b->left; }
const int32_t* or_bvh_right(const OrBvh* b) { return b->right; }
const int32_t* or_bvh_parent(const OrBvh* b) { return b->parent; }
const float* or_bvh_node_boxes(const OrBvh* b) { return b->node_box; }
const float* or_bvh_bounds(const OrBvh* b) { return b->bounds; }
const uint32_t* or_bvh_nodes8(const OrBvh* b) { return b->nodes8; }
const float* or_bvh_prims8(const OrBvh* b) { return b->prims8; }
const int32_t* or_bvh_prim_order8(const OrBvh* b) { return b->prim_order8; }
