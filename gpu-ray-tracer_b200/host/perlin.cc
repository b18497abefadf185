#include "perlin.h"

#include <cmath>
#include <functional>
#include <random>

namespace procedural {

Perlin::Perlin(int seed, int n_sample_vecs) : amplitude_(1.0f), period_(1.0f), n_(n_sample_vecs) {
    std::mt19937 generator{(std::mt19937::result_type)seed};
    auto rng = std::bind(std::uniform_real_distribution<float>{}, generator);  // copies the generator
    sample_vecs_.resize(n_);
    permutation_.resize(n_);
    for (int i = 0; i < n_; i++) {
        float theta = std::acos(2 * rng() - 1);
        float phi = 2 * rng() * M_PI;
        rtb::Float3 v = {std::cos(phi) * std::sin(theta), std::sin(phi) * std::sin(theta), std::cos(theta)};
        sample_vecs_[i] = rtb::normalized(v);
        permutation_[i] = i;
    }
    auto rng_int = std::bind(std::uniform_int_distribution<unsigned>{}, generator);  // seed state again
    for (int i = 0; i < n_; i++) {
        int swap_idx = rng_int() % n_;
        int temp = permutation_[i];
        permutation_[i] = permutation_[swap_idx];
        permutation_[swap_idx] = temp;
    }
}

const rtb::Float3& Perlin::hash(int x, int y, int z) const {
    int hx = x % n_;
    int hxy = (permutation_[hx] + y) % n_;
    int hxyz = (permutation_[hxy] + z) % n_;
    return sample_vecs_[permutation_[hxyz]];
}

static float smoothstep_remap(float d) { return d * d * (3 - 2 * d); }
static float interpolate(float a, float b, float w) { return w * a + (1 - w) * b; }

float Perlin::weight(int ix, int iy, int iz, float mx, float my, float mz, int dx, int dy, int dz) const {
    float cx = ix + dx, cy = iy + dy, cz = iz + dz;
    rtb::Float3 offset = {dx - mx, dy - my, dz - mz};
    const rtb::Float3& wv = hash((int)cx, (int)cy, (int)cz);
    return rtb::dot(wv, rtb::normalized(offset));
}

float Perlin::sample(float x, float y, float z) const {
    float sx = x * n_ / period_, sy = y * n_ / period_, sz = z * n_ / period_;
    int ix = (int)std::floor(sx) % n_, iy = (int)std::floor(sy) % n_, iz = (int)std::floor(sz) % n_;
    float mx = smoothstep_remap(sx - std::floor(sx));
    float my = smoothstep_remap(sy - std::floor(sy));
    float mz = smoothstep_remap(sz - std::floor(sz));
    float w000 = weight(ix, iy, iz, mx, my, mz, 0, 0, 0);
    float w001 = weight(ix, iy, iz, mx, my, mz, 0, 0, 1);
    float w010 = weight(ix, iy, iz, mx, my, mz, 0, 1, 0);
    float w011 = weight(ix, iy, iz, mx, my, mz, 0, 1, 1);
    float w100 = weight(ix, iy, iz, mx, my, mz, 1, 0, 0);
    float w101 = weight(ix, iy, iz, mx, my, mz, 1, 0, 1);
    float w110 = weight(ix, iy, iz, mx, my, mz, 1, 1, 0);
    float w111 = weight(ix, iy, iz, mx, my, mz, 1, 1, 1);
    float x00 = interpolate(w000, w100, mx);
    float x01 = interpolate(w001, w101, mx);
    float x10 = interpolate(w010, w110, mx);
    float x11 = interpolate(w011, w111, mx);
    float xy0 = interpolate(x00, x10, my);
    float xy1 = interpolate(x01, x11, my);
    float xyz = interpolate(xy0, xy1, mz);
    return amplitude_ * xyz;
}

}  // namespace procedural
