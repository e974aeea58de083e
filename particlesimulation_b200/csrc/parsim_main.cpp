// parsim -- drop-in CLI: `parsim <seed> <side> <ncside> <n_part> <n_step>`, same argument
// parsing, output and error behaviour as the reference's main (serial/parsim.cpp:457-489):
// stdout "x0 y0" (3 decimals) then the collision count (:450-453), stderr the wall time of
// the step loop only ("%.1lfs", :475-479), "Error: ..." + exit 1 on failure (:484-488).
// The initial condition is generated on the host exactly like init_particles; the step loop
// runs on the GPU(s) through the C ABI.  PARSIM_NGPU=<n> selects the number of B200s
// (row strips); PARSIM_DEVICE_INIT=1 generates non-negative seeds on the device instead.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iomanip>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/parsim_b200.h"
#include "parsim_host_init.hpp"

namespace {
void check(parsim_ctx *ctx, int rc) {
    if (rc) throw std::runtime_error(parsim_last_error(ctx));
}
}  // namespace

int main(int argc, char *argv[]) {
    parsim_ctx *ctx = nullptr;
    try {
        if (argc != 6) {
            throw std::runtime_error("Usage: " + std::string(argv[0]) +
                                     " <seed> <side_length> <grid_size> <n_particles> <n_timesteps>");
        }
        int seed = std::stol(argv[1]);
        double side_length = std::stod(argv[2]);
        long grid_size = std::stol(argv[3]);
        long long n_particles = std::stoll(argv[4]);
        long n_timesteps = std::stol(argv[5]);
        const char *env = std::getenv("PARSIM_NGPU");
        int n_gpus = env ? std::atoi(env) : 1;
        const char *dev_init = std::getenv("PARSIM_DEVICE_INIT");

        check(nullptr, parsim_create(&ctx, side_length, grid_size, n_particles, n_gpus));
        if (dev_init && std::atoi(dev_init) && seed >= 0) {
            check(ctx, parsim_init_device(ctx, seed));
        } else {
            // stream the generator through a bounded host buffer (8e8 particles = 32 GB otherwise)
            const long long chunk = 1 << 22;
            std::vector<double> buf(5 * (size_t)std::min(chunk, n_particles));
            const size_t stride = buf.size() / 5;
            parsim_host::RandomStream rng(seed);
            check(ctx, parsim_upload_begin(ctx));
            for (long long first = 0; first < n_particles; first += chunk) {
                long long cnt = std::min(chunk, n_particles - first);
                double *x = buf.data(), *y = x + stride, *vx = y + stride, *vy = vx + stride, *m = vy + stride;
                parsim_host::init_particles(rng, side_length, grid_size, n_particles, cnt, x, y, vx, vy, m);
                check(ctx, parsim_upload_chunk(ctx, first, cnt, x, y, vx, vy, m));
            }
            check(ctx, parsim_upload_end(ctx));
        }

        auto t0 = std::chrono::steady_clock::now();
        check(ctx, parsim_run(ctx, n_timesteps));
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "%.1lfs\n", std::chrono::duration<double>(t1 - t0).count());

        double x0 = 0, y0 = 0;
        long long collisions = 0;
        check(ctx, parsim_result(ctx, &x0, &y0, &collisions));
        std::cout << std::fixed << std::setprecision(3) << x0 << " " << y0 << std::endl;
        std::cout << collisions << std::endl;
        parsim_destroy(ctx);
        return 0;
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << std::endl;
        if (ctx) parsim_destroy(ctx);
        return 1;
    }
}
