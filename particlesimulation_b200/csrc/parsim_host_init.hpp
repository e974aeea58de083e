// parsim_host_init.hpp -- host-side (C++) restatement of the reference's seeded initial
// condition, the un-timed part of the drop-in surface: RandomGenerator
// (serial/parsim.cpp:18-48) and ParticleSimulation::init_particles (:220-232).
// Bit-exact evaluation order; five draws per particle from ONE sequential xorshift32 stream.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>

namespace parsim_host {

constexpr double kG = 6.67408e-11;          // :13
constexpr double kEpsilon2 = 0.005 * 0.005; // :14

class RandomStream {
public:
    explicit RandomStream(int input_seed)  // :24
        : state_((uint32_t)std::abs(input_seed) + 987654321u), normal_(input_seed < 0) {}

    double uniform01() {  // :26-32 -- the two states are summed as wrapping 32-bit ints
        uint32_t before = state_;
        state_ ^= state_ << 13;
        state_ ^= state_ >> 17;
        state_ ^= state_ << 5;
        int32_t sum = (int32_t)(before + state_);
        return 0.5 + 0.2328306e-09 * sum;
    }
    double normal01() {  // :34-43
        double r;
        do {
            double u1 = uniform01();
            double u2 = uniform01();
            double z = std::sqrt(-2 * std::log(u1)) * std::cos(2 * M_PI * u2);
            r = 0.5 + 0.15 * z;
        } while (r < 0 || r >= 1);
        return r;
    }
    double next() { return normal_ ? normal01() : uniform01(); }  // :45-47
    uint32_t state() const { return state_; }

private:
    uint32_t state_;
    bool normal_;
};

// particles [first, first+count) of the stream; `rng` must be positioned at particle `first`
inline void init_particles(RandomStream &rng, double side, long ncside, long long n_part,
                           long long count, double *x, double *y, double *vx, double *vy, double *m) {
    for (long long i = 0; i < count; i++) {
        x[i] = rng.next() * side;
        y[i] = rng.next() * side;
        vx[i] = (rng.next() - 0.5) * side / ncside / 5.0;
        vy[i] = (rng.next() - 0.5) * side / ncside / 5.0;
        m[i] = rng.next() * 0.01 * (ncside * ncside) / (size_t)n_part / kG * kEpsilon2;
    }
}

}  // namespace parsim_host
