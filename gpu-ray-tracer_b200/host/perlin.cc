#include "perlin.h"

#include <cmath>
#include <functional>
#include <random>
#include <utility>

namespace procedural {

namespace {

// point on the unit sphere from a polar and an azimuth angle, renormalised like the reference does (perlin.cu:90-92)
rtb::Float3 direction(float polar, float azimuth) {
    const float ring = std::sin(polar);
    const rtb::Float3 v = {std::cos(azimuth) * ring, std::sin(azimuth) * ring, std::cos(polar)};
    return rtb::normalized(v);
}

// the reference's blend puts the weight on its FIRST operand (perlin.cu:8-10)
float blend(float first, float second, float w) { return w * first + (1 - w) * second; }

// smoothstep of the position inside a lattice cell (perlin.cu:35-37)
float fade(float d) { return d * d * (3 - 2 * d); }

struct Axis {  // one coordinate, scaled to lattice units: cell index and faded fraction
    int cell;
    float frac;
    Axis(float scaled, int n) : cell((int)std::floor(scaled) % n), frac(fade(scaled - std::floor(scaled))) {}
};

}  // namespace

Perlin::Perlin(int seed, int n_sample_vecs) : amplitude_(1.0f), period_(1.0f), n_(n_sample_vecs) {
    std::mt19937 engine{(std::mt19937::result_type)seed};
    // std::bind stores a COPY of the engine: the integer stream below starts from the seed state again (perlin.cu:85,96)
    auto uniform01 = std::bind(std::uniform_real_distribution<float>{}, engine);
    gradients_.reserve(n_);
    perm_.reserve(n_);
    for (int k = 0; k < n_; k++) {
        const float polar = std::acos(2 * uniform01() - 1);
        const float azimuth = 2 * uniform01() * M_PI;
        gradients_.push_back(direction(polar, azimuth));
        perm_.push_back(k);
    }
    auto any_unsigned = std::bind(std::uniform_int_distribution<unsigned>{}, engine);
    for (int k = 0; k < n_; k++) {
        const int other = any_unsigned() % n_;
        std::swap(perm_[k], perm_[other]);
    }
}

const rtb::Float3& Perlin::gradient_at(int x, int y, int z) const {
    int h = x % n_;
    h = (perm_[h] + y) % n_;
    h = (perm_[h] + z) % n_;
    return gradients_[perm_[h]];
}

float Perlin::sample(float x, float y, float z) const {
    const Axis ax(x * n_ / period_, n_), ay(y * n_ / period_, n_), az(z * n_ / period_, n_);
    // gradient . unit offset at the eight corners of the cell, corner (dx, dy, dz) at index 4 dx + 2 dy + dz (perlin.cu:49-57)
    float corner[8];
    for (int c = 0; c < 8; c++) {
        const int dx = c >> 2, dy = (c >> 1) & 1, dz = c & 1;
        const rtb::Float3 to_corner = {dx - ax.frac, dy - ay.frac, dz - az.frac};
        corner[c] = rtb::dot(gradient_at(ax.cell + dx, ay.cell + dy, az.cell + dz), rtb::normalized(to_corner));
    }
    // fold the cube along x, then y, then z (perlin.cu:73-80)
    float face[4], edge[2];
    for (int k = 0; k < 4; k++) face[k] = blend(corner[k], corner[4 + k], ax.frac);
    for (int k = 0; k < 2; k++) edge[k] = blend(face[k], face[2 + k], ay.frac);
    return amplitude_ * blend(edge[0], edge[1], az.frac);
}

}  // namespace procedural
