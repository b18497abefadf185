// perlin.h — host-side gradient-noise height field used by the cube-world loader. It has to reproduce the instance
// counts and positions of the reference's worlds, so it follows the behaviour of the reference's procedural::Perlin
// (/root/reference/src/procedural/perlin.cu:8-103) value for value: the gradient table and the permutation come from
// std::mt19937{seed} through libstdc++'s uniform_real_distribution<float> / uniform_int_distribution<unsigned>, and BOTH
// streams start from the seed state (each std::bind copies the generator, perlin.cu:84-85,96); the lattice hash, the
// smoothstep fade and the blend (w * a + (1 - w) * b, weight on the FIRST operand) are the reference's.
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
    // gradient of lattice point (x, y, z): three chained look-ups in the permutation (perlin.cu:12-23)
    const rtb::Float3& gradient_at(int x, int y, int z) const;
    float amplitude_, period_;
    int n_;
    std::vector<rtb::Float3> gradients_;
    std::vector<int> perm_;
};

}  // namespace procedural
#endif
