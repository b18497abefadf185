// rt_exact.cuh — the reference's fp32 arithmetic for the hot path, restated for device code.
//
// Every function below performs the same sequence of correctly rounded fp32 operations
// (+ - * / sqrt, in the reference's order of evaluation) as the reference function it names, so
// that hit decisions, hit distances and colours agree bit for bit with the reference's serial
// path. The translation unit MUST be compiled with -fmad=false (no FMA contraction), default
// -prec-div/-prec-sqrt/-ftz=false. Citations are /root/reference paths.
//
// Used only on CANDIDATES that survived the conservative BVH8 traversal (rt_bvh.cuh), on
// the final hit (normal) and in shading; everything is __host__ __device__ so the host-side unit
// harness (tests/emu) can exercise the same code without a GPU.
#ifndef RTB200_EXACT_CUH
#define RTB200_EXACT_CUH

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define RT_HD __host__ __device__ __forceinline__
// Out-of-line on the device: the exact paths are large (precise sqrt / divide sequences) and are
// reached by few lanes at a time; one copy per kernel keeps the traversal loop inside the
// instruction cache (k_shade was 199 KB of SASS with everything inlined, stall_no_instruction 8.0).
#ifdef __CUDACC__
#define RT_HD_NOINLINE inline __host__ __device__ __noinline__
#else
#define RT_HD_NOINLINE inline
#endif

// include/raymath/linear.h:15 — `constexpr double THRESHOLD = 1E-5f`: an fp32 value widened to
// double; comparing an fp32 operand against it is the same as comparing against the fp32 constant.
#define RT_THR 1E-5f

namespace rtx {

struct Ray { float3 o, d; };
struct Quat { float i, j, k, r; };
struct Pose { Quat q; float3 p; int identity_rot; };

RT_HD float3 mk3(float x, float y, float z) { float3 v; v.x = x; v.y = y; v.z = z; return v; }
RT_HD float3 add3(float3 a, float3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD float3 sub3(float3 a, float3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
// rmath::operator*(T, Vec) multiplies each coordinate by the coefficient (linear.h:127-133)
RT_HD float3 scale3(float c, float3 v) { return mk3(v.x * c, v.y * c, v.z * c); }
RT_HD float3 neg3(float3 v) { return scale3(-1.0f, v); }
// linear.h:197-204: the sum starts at 0 and accumulates in index order
RT_HD float dot3(float3 a, float3 b) {
    float s = 0.0f;
    s += a.x * b.x;
    s += a.y * b.y;
    s += a.z * b.z;
    return s;
}
RT_HD float len3(float3 v) { return sqrtf(dot3(v, v)); }
// linear.h:160-168: zero vector when the length is not above the threshold
RT_HD float3 normalized3(float3 v) {
    float len = len3(v);
    if (len > RT_THR) return scale3(1.0f / len, v);
    return mk3(0.0f, 0.0f, 0.0f);
}
// linear.h:206-213
RT_HD float3 cross3(float3 a, float3 b) {
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

RT_HD float4 mk4(float x, float y, float z, float w) { float4 v; v.x = x; v.y = y; v.z = z; v.w = w; return v; }
RT_HD float4 add4(float4 a, float4 b) { return mk4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
RT_HD float4 mul4(float4 a, float4 b) { return mk4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
RT_HD float4 scale4(float c, float4 a) { return mk4(a.x * c, a.y * c, a.z * c, a.w * c); }
RT_HD float4 ld4(const float* p) { return mk4(p[0], p[1], p[2], p[3]); }

// ---- rmath::Quat (include/raymath/geometry.h:17-204) ----
RT_HD float qdot(Quat a, Quat b) {
    float s = 0.0f;
    s += a.i * b.i;
    s += a.j * b.j;
    s += a.k * b.k;
    s += a.r * b.r;
    return s;
}
RT_HD Quat qnormalized(Quat q) {  // geometry.h:158-161
    float len = sqrtf(qdot(q, q));
    Quat z = {0.0f, 0.0f, 0.0f, 0.0f};
    if (len > RT_THR) {
        float inv = 1.0f / len;
        Quat r = {q.i * inv, q.j * inv, q.k * inv, q.r * inv};
        return r;
    }
    return z;
}
RT_HD Quat qinverse(Quat q) {  // geometry.h:145-156
    float sq = qdot(q, q);
    Quat z = {0.0f, 0.0f, 0.0f, 0.0f};
    if (sq < RT_THR) return z;
    float inv = 1.0f / sq;
    Quat r = {q.i * -inv, q.j * -inv, q.k * -inv, q.r * inv};
    return r;
}
RT_HD Quat qmul(Quat a, Quat o) {  // geometry.h:163-170 (Hamilton product, this evaluation order)
    Quat r;
    r.i = a.i * o.r + a.r * o.i + a.j * o.k - a.k * o.j;
    r.j = a.j * o.r + a.r * o.j + a.k * o.i - a.i * o.k;
    r.k = a.k * o.r + a.r * o.k + a.i * o.j - a.j * o.i;
    r.r = a.r * o.r - a.i * o.i - a.j * o.j - a.k * o.k;
    return r;
}
// geometry.h:176-181 operator*(Quat, Vec3): sandwich product, then the input length restored.
RT_HD_NOINLINE float3 qrotate(Quat q, float3 v) {
    float length = len3(v);
    Quat qv = {v.x, v.y, v.z, 0.0f};
    Quat prod = qmul(qmul(qnormalized(q), qv), qinverse(q));
    return scale3(length, normalized3(mk3(prod.i, prod.j, prod.k)));
}
// The same operator for q == (0,0,0,1) exactly: both quaternion products reproduce v bit for bit
// (every term is v*1 or +-0), so what remains is normalise-then-rescale. Every shipped world has
// identity rotations on all instances and meshes (src/procedural/cube_world.cc:160-164).
RT_HD float3 qrotate_identity(float3 v) {
    float length = len3(v);
    return scale3(length, normalized3(v));
}

// ---- rprimitives::Entity (src/rayprimitives/entity.cu:5-37) ----
RT_HD float3 vec_to_local(const Pose& e, float3 v) { return e.identity_rot ? qrotate_identity(v) : qrotate(e.q, v); }
RT_HD float3 point_to_local(const Pose& e, float3 v) { return vec_to_local(e, sub3(v, e.p)); }
RT_HD float3 vec_from_local(const Pose& e, float3 v) {
    return e.identity_rot ? qrotate_identity(v) : qrotate(qinverse(e.q), v);
}
RT_HD float3 point_from_local(const Pose& e, float3 v) { return add3(vec_from_local(e, v), e.p); }

// geometry.h:216: the Ray constructor normalises the direction
RT_HD Ray ray_make(float3 o, float3 d) { Ray r; r.o = o; r.d = normalized3(d); return r; }
RT_HD float3 ray_at(const Ray& r, float t) { return add3(r.o, scale3(t, r.d)); }

// ---- ropt::BoundingBox::intersects (src/rayopt/bounding_box.cu:62-104) ----
// Kay/Kajiya slabs with real divides; an axis with dir == 0 is skipped; no test against the
// current closest hit. box = {min.xyz, max.xyz}; degenerate boxes never hit.
RT_HD bool ref_box_intersects(const float* box, bool nondegenerate, const Ray& r) {
    if (!nondegenerate) return false;
    const float dir[3] = {r.d.x, r.d.y, r.d.z};
    const float org[3] = {r.o.x, r.o.y, r.o.z};
    float time_min = -INFINITY, time_max = INFINITY;
#pragma unroll
    for (int axis = 0; axis < 3; axis++) {
        if (dir[axis] == 0.0f) continue;
        float time_near = (box[axis] - org[axis]) / dir[axis];
        float time_far = (box[3 + axis] - org[axis]) / dir[axis];
        if (time_near > time_far) { float tmp = time_near; time_near = time_far; time_far = tmp; }
        if (time_near > time_min) time_min = time_near;
        if (time_far < time_max) time_max = time_far;
        if (time_min > time_max || time_max < RT_THR) return false;
    }
    return true;
}

// ---- rmath::Plane::hit + rmath::Triangle::hit (include/raymath/geometry.h:253-261, 275-290) ----
// Unsigned-area barycentric test: accept iff |b0+b1+b2-1| <= THRESHOLD. uv = (b1, b2).
// One unsigned sub-area over the triangle's area (geometry.h:283-285). Out of line on the device: three calls instead
// of three inlined copies of cross + sqrt + divide keep the exact-test code small (instruction cache, rt_engine.cu k_tile).
RT_HD_NOINLINE float ref_bary_term(float3 x, float3 y, float3 p, float tri_area) {
    return len3(cross3(sub3(x, p), sub3(y, p))) / tri_area;
}
RT_HD bool ref_triangle_hit(float3 a, float3 b, float3 c, const Ray& r, float& time, float& u, float& v) {
    float3 plane_norm = cross3(sub3(b, a), sub3(c, a));
    float3 n = normalized3(plane_norm);
    float denom = dot3(r.d, n);
    if (fabsf(denom) < RT_THR) return false;
    time = (1.0f / denom) * dot3(sub3(a, r.o), n);
    float3 p = ray_at(r, time);
    float tri_area = len3(plane_norm);
    float bary0 = ref_bary_term(c, b, p, tri_area);
    float bary1 = ref_bary_term(c, a, p, tri_area);
    float bary2 = ref_bary_term(a, b, p, tri_area);
    if (fabsf(bary0 + bary1 + bary2 - 1.0f) <= RT_THR) {
        u = bary1;
        v = bary2;
        return true;
    }
    return false;
}

// The ray-independent part of rmath::Triangle::hit: plane normal (geometry.h:276), its normalisation (Plane ctor) and the
// triangle's area (geometry.h:282). Computed once per mesh triangle at scene set-up with the very functions
// ref_triangle_hit uses, so ref_triangle_hit_pre returns the same bits with two square roots and a divide fewer per candidate.
struct TriPre { float3 n; float area; };
RT_HD TriPre ref_triangle_pre(float3 a, float3 b, float3 c) {
    float3 plane_norm = cross3(sub3(b, a), sub3(c, a));
    TriPre p;
    p.n = normalized3(plane_norm);
    p.area = len3(plane_norm);
    return p;
}
RT_HD bool ref_triangle_hit_pre(float3 a, float3 b, float3 c, float3 n, float tri_area, const Ray& r, float& time, float& u, float& v) {
    float denom = dot3(r.d, n);
    if (fabsf(denom) < RT_THR) return false;
    time = (1.0f / denom) * dot3(sub3(a, r.o), n);
    float3 p = ray_at(r, time);
    float bary0 = ref_bary_term(c, b, p, tri_area);
    float bary1 = ref_bary_term(c, a, p, tri_area);
    float bary2 = ref_bary_term(a, b, p, tri_area);
    if (fabsf(bary0 + bary1 + bary2 - 1.0f) <= RT_THR) {
        u = bary1;
        v = bary2;
        return true;
    }
    return false;
}

// ---- rmath::reflect / refract (include/raymath/linear.h:215-242) ----
RT_HD float3 ref_reflect(float3 dir, float3 norm) {
    float d_len = len3(dir);
    float3 d_norm = normalized3(dir);
    float3 n_norm = normalized3(norm);
    float3 projection = scale3(dot3(d_norm, n_norm), n_norm);
    float3 r_norm = normalized3(sub3(d_norm, scale3(2.0f, projection)));
    return scale3(d_len, r_norm);
}
RT_HD float3 ref_refract(float3 dir, float3 norm, float index_from, float index_to, bool& tir) {
    float d_len = len3(dir);
    float3 d_norm = normalized3(dir);
    float3 n_norm = normalized3(norm);
    float index_ratio = index_from / index_to;
    float cosi = dot3(d_norm, n_norm);
    float sint_2 = index_ratio * index_ratio * (1.0f - cosi * cosi);
    if (sint_2 > 1.0f) {
        tir = true;
        return scale3(d_len, ref_reflect(d_norm, n_norm));
    }
    tir = false;
    return scale3(d_len, add3(scale3(index_ratio, d_norm), scale3(index_ratio * cosi - sqrtf(1.0f - sint_2), n_norm)));
}

// powf as the reference's host build evaluates it (glibc: correctly rounded in all but a
// vanishing fraction of cases): a double-precision pow rounded once to fp32. Only reached for
// refractive materials (trans_atten / calc_shadow_atten) and specular terms; pow(x, 0) == 1.
RT_HD_NOINLINE float ref_powf(float x, float y) {
    if (y == 0.0f) return 1.0f;
    if (y == 1.0f) return x;
    return (float)pow((double)x, (double)y);
}

// ---- renv::Camera::at + Quat::to_Mat3 (src/rayenv/camera.cu:32-42, geometry.h:183-198) ----
struct CameraBasis { float3 r, u, f, pos; float near_, unit, half_w, half_h; };

RT_HD CameraBasis camera_basis(const float* rot, const float* pos, float global_near, float unit_to_pixels,
                               float canvas_w, float canvas_h) {
    Quat q = {rot[0], rot[1], rot[2], rot[3]};
    float length = sqrtf(qdot(q, q));
    float ni = q.i / length, nj = q.j / length, nk = q.k / length, nr = q.r / length;
    float ii = 2.0f * ni * ni, jj = 2.0f * nj * nj, kk = 2.0f * nk * nk;
    float ri = 2.0f * nr * ni, rj = 2.0f * nr * nj, rk = 2.0f * nr * nk;
    float ij = 2.0f * ni * nj, ik = 2.0f * ni * nk, jk = 2.0f * nj * nk;
    CameraBasis c;
    c.r = normalized3(mk3(1.0f - (jj + kk), ij + rk, ik - rj));
    c.u = normalized3(mk3(ij - rk, 1.0f - (ii + kk), jk + ri));
    c.f = normalized3(mk3(ik + rj, jk - ri, 1.0f - (ii + jj)));
    c.pos = mk3(pos[0], pos[1], pos[2]);
    c.near_ = global_near;
    c.unit = unit_to_pixels;
    c.half_w = 0.5f * canvas_w;
    c.half_h = 0.5f * canvas_h;
    return c;
}
// The basis depends only on the camera, so it is computed once per frame; per pixel only the
// remaining operations of Camera::at run, in the reference's order.
RT_HD Ray camera_ray(const CameraBasis& c, float canvas_x, float canvas_y) {
    float gx = (canvas_x - c.half_w) / c.unit;
    float gy = (c.half_h - canvas_y) / c.unit;
    float3 dir = add3(add3(scale3(c.near_, c.f), scale3(gx, c.r)), scale3(gy, c.u));
    return ray_make(c.pos, dir);
}

// ---- src/raytracer.cu:37-40 clamp, include/rayenv/color.h:41-42 truncation, color.cu:24-26 ----
RT_HD uint32_t pack_color(float4 c) {
    float ch[4] = {c.x, c.y, c.z, c.w};
    uint32_t out[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float v = ch[i] > 1.0f ? 1.0f : ch[i];
        out[i] = (uint32_t)(uint8_t)(int)(255.0f * v);
    }
    return (uint32_t)((int)(out[0] << 24) + (int)(out[1] << 16) + (int)(out[2] << 8) + (int)out[3]);
}

}  // namespace rtx
#endif
