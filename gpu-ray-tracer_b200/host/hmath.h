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

}  // namespace rtb

#endif
