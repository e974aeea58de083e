// parsim -- drop-in CLI: `parsim <seed> <side> <ncside> <n_part> <n_step>`, same argument
// parsing, output and error behaviour as the reference's main (serial/parsim.cpp:457-489):
// stdout "x0 y0" (3 decimals) then the collision count (:450-453), stderr the wall time of
// the step loop only ("%.1lfs", :475-479), "Error: ..." + exit 1 on failure (:484-488).
// The initial condition is init_particles': for non-negative seeds generated on the device (bit-
// identical to the host generator; PARSIM_DEVICE_INIT=0 streams it from the host instead), for
// negative seeds (the clustered case) on the host; the step loop runs on the GPU(s) through the
// C ABI.  PARSIM_NGPU=<n> selects the number of B200s (row strips);
// PARSIM_BALANCE=1 cuts the strips so that they hold equal numbers of PARTICLES (clustered
// inputs, i.e. negative seeds) instead of equal numbers of rows.  PARSIM_TRACE=<file> writes the
// per-particle trace the reference keeps commented out in simulate() (serial/parsim.cpp:424-431,
// :442-446: all particles and cells before the loop, then "t=<i>" and all particles after every
// step) for diffing against the oracle; the five positionals and stdout stay untouched.
// PARSIM_SEMANTICS=omp selects the OpenMP variant's list update (no strays).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
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

// "Particle j: mass=... x=... y=... vx=... vy=..." for every particle (:426, :445); dead
// particles keep their frozen state with mass 0, like the reference's tombstones
void trace_particles(parsim_ctx *ctx, long long n, std::ofstream &out) {
    std::vector<double> x(n), y(n), vx(n), vy(n), m(n);
    std::vector<int> alive(n);
    check(ctx, parsim_dump(ctx, x.data(), y.data(), vx.data(), vy.data(), m.data(), alive.data(), nullptr,
                           nullptr, nullptr, nullptr));
    out << std::fixed << std::setprecision(3);
    for (long long j = 0; j < n; j++)
        out << "Particle " << j << ": mass=" << (alive[j] ? m[j] : 0.0) << " x=" << x[j] << " y=" << y[j]
            << " vx=" << vx[j] << " vy=" << vy[j] << "\n";
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

        const char *balance = std::getenv("PARSIM_BALANCE");
        const long long chunk_len = 1 << 22;
        if (n_gpus > 1 && balance && std::atoi(balance)) {
            // first pass over the seeded stream: particles per grid row -> strip boundaries
            std::vector<long long> rows((size_t)grid_size, 0);
            std::vector<long> first((size_t)n_gpus + 1);
            std::vector<double> buf(5 * (size_t)std::min(chunk_len, n_particles));
            const size_t stride = buf.size() / 5;
            parsim_host::RandomStream rng(seed);
            for (long long at = 0; at < n_particles; at += chunk_len) {
                long long cnt = std::min(chunk_len, n_particles - at);
                double *x = buf.data(), *y = x + stride, *vx = y + stride, *vy = vx + stride, *m = vy + stride;
                parsim_host::init_particles(rng, side_length, grid_size, n_particles, cnt, x, y, vx, vy, m);
                check(nullptr, parsim_host_row_counts(side_length, grid_size, cnt, y, rows.data()));
            }
            check(nullptr, parsim_balance_rows(grid_size, n_gpus, rows.data(), first.data()));
            check(nullptr, parsim_create_rows(&ctx, side_length, grid_size, n_particles, n_gpus, first.data()));
        } else {
            check(nullptr, parsim_create(&ctx, side_length, grid_size, n_particles, n_gpus));
        }
        // non-negative seeds: the uniform generator runs on the device (xorshift32 jump-ahead, bit-identical
        // to the host generator: tests/test_gpu_parity.py) unless PARSIM_DEVICE_INIT=0 asks for the host stream
        if (seed >= 0 && !(dev_init && std::atoi(dev_init) == 0)) {
            check(ctx, parsim_init_device(ctx, seed));
        } else {
            // stream the generator through a bounded host buffer (8e8 particles = 32 GB otherwise)
            const long long chunk = chunk_len;
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

        if (const char *sem = std::getenv("PARSIM_SEMANTICS")) {  // "serial" (default) | "omp"
            const std::string v(sem);
            if (v == "omp") check(ctx, parsim_set_option(ctx, "semantics", 1));
            else if (v != "serial" && !v.empty()) throw std::runtime_error("PARSIM_SEMANTICS must be serial or omp");
        }
        const char *trace = std::getenv("PARSIM_TRACE");
        std::ofstream tout;
        if (trace && *trace) {
            if (n_gpus > 1) throw std::runtime_error("PARSIM_TRACE needs a single GPU (PARSIM_NGPU=1)");
            tout.open(trace);
            if (!tout) throw std::runtime_error(std::string("cannot open trace file ") + trace);
            trace_particles(ctx, n_particles, tout);
            tout << std::fixed << std::setprecision(3);
            for (long long j = 0; j < (long long)grid_size * grid_size; j++)  // cells are still zero here (:215)
                tout << "Cell " << j << " x: " << 0.0 << " y: " << 0.0 << " m: " << 0.0 << "\n";
        }
        auto t0 = std::chrono::steady_clock::now();
        if (tout.is_open()) {
            for (long i = 0; i < n_timesteps; i++) {
                check(ctx, parsim_run(ctx, 1));
                tout << "t=" << i << "\n";
                trace_particles(ctx, n_particles, tout);
            }
            tout.close();
        } else {
            check(ctx, parsim_run(ctx, n_timesteps));
        }
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
