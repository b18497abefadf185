// hmath.h — host-side fp32 helpers for scene construction.
//
// These run once at scene-load time (never per ray). They follow the evaluation order of the
// reference's rmath templates (/root/reference/include/raymath/linear.h:24-213,
// include/raymath/geometry.h:34-41,183-198) so that the flattened scene handed to the GPU is
// bit-identical to what the reference's SceneBuilder produces. Build with -ffp-contract=off.
#ifndef RTB200_HMATH_H
#define RTB200_HMATH_H

#include <cmath>

namespace rtb {

constexpr float kThreshold = 1E-5f;  // rmath::THRESHOLD (linear.h:15), exactly representable in fp32

struct Float3 {
    float x, y, z;
};
struct Float4 {
    float x, y, z, w;
};
// quaternion stored (i, j, k, r) like rmath::Quat::inner
struct Quat {
    float i, j, k, r;
    static Quat identity() { return Quat{0.f, 0.f, 0.f, 1.f}; }
};

inline Float3 operator+(Float3 a, Float3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Float3 operator-(Float3 a, Float3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Float3 operator*(float c, Float3 v) { return {v.x * c, v.y * c, v.z * c}; }
inline Float4 operator*(float c, Float4 v) { return {v.x * c, v.y * c, v.z * c, v.w * c}; }

inline float dot(Float3 a, Float3 b) {
    float s = 0;
    s += a.x * b.x;
    s += a.y * b.y;
    s += a.z * b.z;
    return s;
}
inline float length(Float3 v) { return std::sqrt(dot(v, v)); }
inline Float3 normalized(Float3 v) {
    float len = length(v);
    if (len > kThreshold) return (1 / len) * v;
    return {0.f, 0.f, 0.f};
}
inline Float3 cross(Float3 a, Float3 b) {
    float x = a.y * b.z - a.z * b.y;
    float y = a.z * b.x - a.x * b.z;
    float z = a.x * b.y - a.y * b.x;
    return {x, y, z};
}
// rmath::Quat(axis, theta) (geometry.h:34-41): theta in RADIANS, no normalisation of the axis
inline Quat quat_axis_angle(Float3 axis, float theta) {
    float half_cos_theta = std::cos(0.5f * theta);
    float half_sin_theta = std::sin(0.5f * theta);
    return Quat{axis.x * half_sin_theta, axis.y * half_sin_theta, axis.z * half_sin_theta, half_cos_theta};
}
// Hamilton product in the reference's evaluation order (geometry.h:163-170)
inline Quat operator*(Quat a, Quat o) {
    Quat r;
    r.i = a.i * o.r + a.r * o.i + a.j * o.k - a.k * o.j;
    r.j = a.j * o.r + a.r * o.j + a.k * o.i - a.i * o.k;
    r.k = a.k * o.r + a.r * o.k + a.i * o.j - a.j * o.i;
    r.r = a.r * o.r - a.i * o.i - a.j * o.j - a.k * o.k;
    return r;
}

// ---- motion: the operations behind Entity::translate / rotate and the mouse-look
// of the reference's event loop (include/rayprimitives/entity.h:49-76, src/main.cc:140-179) ----
inline float qdot(Quat a, Quat b) {
    float s = 0;
    s += a.i * b.i;
    s += a.j * b.j;
    s += a.k * b.k;
    s += a.r * b.r;
    return s;
}
// Quat::normalized (geometry.h:158-161) = Vec4::normalized (linear.h:160-168)
inline Quat qnormalized(Quat q) {
    float len = std::sqrt(qdot(q, q));
    if (len > kThreshold) {
        float inv = 1 / len;
        return Quat{q.i * inv, q.j * inv, q.k * inv, q.r * inv};
    }
    return Quat{0.f, 0.f, 0.f, 0.f};
}
// Quat::inverse (geometry.h:145-156)
inline Quat qinverse(Quat q) {
    float sq = qdot(q, q);
    if (sq < kThreshold) return Quat{0.f, 0.f, 0.f, 0.f};
    float inv = 1 / sq;
    return Quat{q.i * -inv, q.j * -inv, q.k * -inv, q.r * inv};
}
// operator*(Quat, Vec3) (geometry.h:176-181): sandwich product, input length restored
inline Float3 operator*(Quat q, Float3 v) {
    float len = length(v);
    Quat prod = qnormalized(q) * Quat{v.x, v.y, v.z, 0.f} * qinverse(q);
    return len * normalized(Float3{prod.i, prod.j, prod.k});
}
// columns of Quat::to_Mat3 (geometry.h:183-198), normalised twice like Camera::right/up/forward
// followed by the Ray constructor do (src/rayenv/camera.cu:11-30, geometry.h:216)
struct Basis { Float3 right, up, forward; };
inline Basis quat_basis(Quat q) {
    float len = std::sqrt(qdot(q, q));
    float ni = q.i / len, nj = q.j / len, nk = q.k / len, nr = q.r / len;
    float ii = 2.0f * ni * ni, jj = 2.0f * nj * nj, kk = 2.0f * nk * nk;
    float ri = 2.0f * nr * ni, rj = 2.0f * nr * nj, rk = 2.0f * nr * nk, ij = 2.0f * ni * nj, ik = 2.0f * ni * nk,
          jk = 2.0f * nj * nk;
    Basis b;
    b.right = normalized(normalized(Float3{1 - (jj + kk), ij + rk, ik - rj}));
    b.up = normalized(normalized(Float3{ij - rk, 1 - (ii + kk), jk + ri}));
    b.forward = normalized(normalized(Float3{ik + rj, jk - ri, 1 - (ii + jj)}));
    return b;
}

}  // namespace rtb

#endif
