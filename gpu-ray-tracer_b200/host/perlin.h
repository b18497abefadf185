// perlin.h — host-side Perlin height field used by the cube-world loader. Mirrors the behaviour
// of the reference's procedural::Perlin (/root/reference/src/procedural/perlin.cu:8-103): the
// gradient table and permutation come from std::mt19937{seed} through libstdc++'s
// uniform_real_distribution<float> / uniform_int_distribution<unsigned>, and BOTH streams start
// from the seed state (each std::bind copies the generator, perlin.cu:84-85,96).
#ifndef RTB200_PERLIN_H
#define RTB200_PERLIN_H

#include <vector>

#include "hmath.h"

namespace procedural {

class Perlin {
public:
    Perlin(int seed, int n_sample_vecs);
    void set_amplitude(float a) { amplitude_ = a; }
    void set_period(float p) { period_ = p; }
    float sample(float x, float y, float z) const;
    int n_sample_vecs() const { return n_; }

private:
    const rtb::Float3& hash(int x, int y, int z) const;
    float weight(int ix, int iy, int iz, float mx, float my, float mz, int dx, int dy, int dz) const;
    float amplitude_, period_;
    int n_;
    std::vector<rtb::Float3> sample_vecs_;
    std::vector<int> permutation_;
};

}  // namespace procedural
#endif
