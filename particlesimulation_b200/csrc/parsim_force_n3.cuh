// parsim_force_n3.cuh -- force + integration with Newton's third law inside the cell (updateForces +
// applyForce, serial/parsim.cpp:292-377,:150-195).
//
// The reference sums every in-cell pair ONCE (j < k, :356-366) and applies the reaction to k
// (:143-147).  k_force_integrate (parsim_kernels.cuh) evaluates ORDERED pairs instead -- 13 FP64
// instructions per ordered pair, 26 per unordered pair -- because a thread owns one target.  Here a
// WARP owns a cell and keeps the cell's particles in registers, K per lane ("chunks" of m = ceil(n/K)
// consecutive list entries; lane l holds entries l, m + l, 2m + l, ...).  Partners are fetched by
// warp shuffle at a lane offset s and the reaction is shuffled back:
//
//   step 0          : pairs of the lane's own K particles (no shuffle)
//   step s = 1..(m-1)/2 : every lane fetches the particle of lane (l + s) mod m of each chunk, interacts
//                     it with its own K particles, and returns the summed reaction to lane (l + s) mod m
//   step m/2 (m even): the same with lanes l < m/2 only (each pair is met from both ends otherwise)
//
// Every unordered pair {(c, a), (c', b)} is met exactly once: at s = (b - a) mod m if that is below
// m/2, else from the other end.  16 FP64 instructions + 1 MUFU per UNORDERED pair (the two masses
// make two weights) against 26, at the price of 10 shuffles per K pairs and idle lanes when n is not
// a multiple of 32 K.  The order of the additions into a particle's accumulator is a function of the
// cell's list (n and the particle's place in it) only: no dependence on scheduling, block shape or
// the number of GPUs, so N strips still equal one GPU bit for bit.  It is NOT the summation order
// of k_force_integrate: accelerations differ in the last bits (as both differ from the reference's,
// SURVEY 0 fact 5), which is why this kernel is a separate `semantics`-neutral variant with its own
// oracle tolerance tests and is never mixed with the ordered kernels inside one run.
#pragma once
#include "parsim_kernels.cuh"
#include "parsim_force_pairs2.cuh"  // integrate_store

namespace parsim {

// one unordered pair: `o` = the lane's own particle, `f` = the fetched one; d = f - o.
// a_o += m_f d / |d|^3 ; the reaction on f is accumulated as r += m_o d / |d|^3 (the caller subtracts it)
__device__ __forceinline__ void pair_n3(double dx, double dy, double m_o, double m_f, double &ax, double &ay,
                                        double &rx, double &ry) {
    const double d2 = fma(dx, dx, fma(dy, dy, kTinyD2));
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d2));
    const double yy = y * y;
    const double e = fma(-d2, yy, 1.0);
    const double p = fma(e, 1.875, 1.5);
    const double y3 = y * yy;
    const double r3 = fma(y3 * e, p, y3);  // d2^(-3/2)
    const double wo = m_f * r3, wf = m_o * r3;
    ax = fma(wo, dx, ax);
    ay = fma(wo, dy, ay);
    rx = fma(wf, dx, rx);
    ry = fma(wf, dy, ry);
}

struct ForceArgs {
    const double2 *pos_in;
    const double *m;
    const int *cell_start;
    const double2 *com_xy;
    const double *com_m;
    double2 *pos_out;
    float2 *posf_out;
    double2 *vel;
    int *newcell;
    unsigned char *flags;
    int *mover_list;
    int *n_movers;
    double2 *dbg_acc;
};

// The 8 neighbour centres of mass of local cell c, shifted across the periodic border, in the
// reference's order (dx outer, dy inner, :301-305): lane j < 8 loads entry j; the callers broadcast
// them one at a time (three doubles live instead of twenty-four).
struct Neighbour { double x, y, m; };
__device__ __forceinline__ Neighbour neighbour_of_cell(const Geom &g, int c, const ForceArgs &a, int lane) {
    Neighbour v{0.0, 0.0, 0.0};
    if (lane < 8) {
        const int j = lane;
        const int dxn = j < 3 ? -1 : j < 5 ? 0 : 1;
        const int dyn = j < 3 ? j - 1 : j == 3 ? -1 : j == 4 ? 1 : j - 6;
        const int lrow = c / g.nc;
        int gx = c - lrow * g.nc + dxn;
        double sx = 0.0, sy = 0.0;
        if (gx >= g.nc) { sx = g.side; gx -= g.nc; } else if (gx < 0) { sx = -g.side; gx += g.nc; }
        const int gy = g.row0 + lrow + dyn;
        if (gy >= g.nc) sy = g.side; else if (gy < 0) sy = -g.side;
        const int q = (lrow + dyn + 1) * g.nc + gx;
        const double2 cm = __ldg(&a.com_xy[q]);
        v.x = sx + cm.x;
        v.y = sy + cm.y;
        v.m = __ldg(&a.com_m[q]);
    }
    return v;
}

// a cell of n <= 32 K particles, list entries [s, s + n), handled by one warp
template <int K>
__device__ __forceinline__ void n3_cell(const Geom &g, int c, int s, int n, const ForceArgs &a, int lane) {
    const int m = (n + K - 1) / K;  // lanes in use; chunk k holds entries k m .. k m + m - 1
    double x[K], y[K], mk[K], ax[K], ay[K];
    bool ok[K];
#pragma unroll
    for (int k = 0; k < K; k++) {
        const int idx = k * m + lane;
        ok[k] = lane < m && idx < n;
        double2 p = make_double2(0.0, 0.0);
        double mm = 0.0;  // a massless particle at the origin: every weight it is part of is exactly 0
        if (ok[k]) {
            p = a.pos_in[s + idx];
            mm = a.m[s + idx];
        }
        x[k] = p.x; y[k] = p.y; mk[k] = mm;
        ax[k] = 0.0; ay[k] = 0.0;
    }
    {   // the 8 neighbour centres of mass (:369-374); the own cell's centre of mass is never a source (:305)
        const Neighbour v = neighbour_of_cell(g, c, a, lane);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const double nx = __shfl_sync(0xffffffffu, v.x, j), ny = __shfl_sync(0xffffffffu, v.y, j);
            const double nm = __shfl_sync(0xffffffffu, v.m, j);
#pragma unroll
            for (int k = 0; k < K; k++) accumulate(nx - x[k], ny - y[k], nm, ax[k], ay[k]);
        }
    }
    // step 0: the lane's own particles among themselves
#pragma unroll
    for (int k = 0; k < K; k++)
#pragma unroll
        for (int q = k + 1; q < K; q++) {
            double rx = 0.0, ry = 0.0;
            pair_n3(x[q] - x[k], y[q] - y[k], mk[k], mk[q], ax[k], ay[k], rx, ry);
            ax[q] -= rx;
            ay[q] -= ry;
        }
    const int full = (m - 1) >> 1;
#pragma unroll 1
    for (int st = 1; st <= (m >> 1); st++) {
        int src = lane + st;
        src = src >= m ? src - m : src;
        int dst = lane - st;
        dst = dst < 0 ? dst + m : dst;
        const bool mine = st <= full || lane < (m >> 1);  // the half step of an even m: one end only
#pragma unroll
        for (int f = 0; f < K; f++) {
            const double fx = __shfl_sync(0xffffffffu, x[f], src);
            const double fy = __shfl_sync(0xffffffffu, y[f], src);
            double fm = __shfl_sync(0xffffffffu, mk[f], src);
            fm = mine ? fm : 0.0;
            double rx = 0.0, ry = 0.0;
#pragma unroll
            for (int k = 0; k < K; k++)
                pair_n3(fx - x[k], fy - y[k], mine ? mk[k] : 0.0, fm, ax[k], ay[k], rx, ry);
            ax[f] -= __shfl_sync(0xffffffffu, rx, dst);
            ay[f] -= __shfl_sync(0xffffffffu, ry, dst);
        }
    }
#pragma unroll
    for (int k = 0; k < K; k++) {
        if (!ok[k]) continue;
        const int i = s + k * m + lane;
        integrate_store(g, i, c, make_double2(x[k], y[k]), a.vel[i], ax[k], ay[k], a.pos_out, a.posf_out, a.vel,
                        a.newcell, a.flags, a.mover_list, a.n_movers, a.dbg_acc);
    }
}

// cells above the register capacity of the instantiation: ordered sweep, sources through L1.  The
// summation order is again a function of the list only (neighbours, then the list in order).
__device__ __noinline__ void n3_cell_large(const Geom &g, int c, int s, int n, const ForceArgs &a, int lane) {
    const Neighbour v = neighbour_of_cell(g, c, a, lane);
    for (int t0 = 0; t0 < n; t0 += 32) {  // whole warp iterates: the broadcasts need every lane
        const int t = t0 + lane;
        const bool ok = t < n;
        const int i = s + (ok ? t : 0);
        const double2 p = a.pos_in[i];
        double ax = 0.0, ay = 0.0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const double nx = __shfl_sync(0xffffffffu, v.x, j), ny = __shfl_sync(0xffffffffu, v.y, j);
            accumulate(nx - p.x, ny - p.y, __shfl_sync(0xffffffffu, v.m, j), ax, ay);
        }
#pragma unroll 4
        for (int k = s; k < s + n; k++) {
            const double2 q = __ldg(&a.pos_in[k]);
            accumulate(q.x - p.x, q.y - p.y, __ldg(&a.m[k]), ax, ay);
        }
        if (!ok) continue;
        integrate_store(g, i, c, p, a.vel[i], ax, ay, a.pos_out, a.posf_out, a.vel, a.newcell, a.flags,
                        a.mover_list, a.n_movers, a.dbg_acc);
    }
}

constexpr int kN3Warps = 4;  // cells per block

// KMAX = particles per lane the instantiation has registers for (cells up to 32 KMAX entries)
template <int KMAX, int MINB>
__global__ void __launch_bounds__(kN3Warps * 32, MINB) k_force_n3(Geom g, ForceArgs a) {
    const int lane = threadIdx.x & 31;
    const int c = blockIdx.x * kN3Warps + (threadIdx.x >> 5);
    if (c >= g.ncells) return;  // warp-uniform
    const int s = a.cell_start[c], n = a.cell_start[c + 1] - s;
    if (n == 0) return;
    const int K = (n + 31) >> 5;
    if (K > KMAX) { n3_cell_large(g, c, s, n, a, lane); return; }
    switch (K) {  // warp-uniform
    case 1: n3_cell<1>(g, c, s, n, a, lane); break;
    case 2: if (KMAX >= 2) n3_cell<2>(g, c, s, n, a, lane); break;
    case 3: if (KMAX >= 3) n3_cell<3>(g, c, s, n, a, lane); break;
    case 4: if (KMAX >= 4) n3_cell<4>(g, c, s, n, a, lane); break;
    case 5: if (KMAX >= 5) n3_cell<5>(g, c, s, n, a, lane); break;
    case 6: if (KMAX >= 6) n3_cell<6>(g, c, s, n, a, lane); break;
    default: n3_cell_large(g, c, s, n, a, lane); break;
    }
}

}  // namespace parsim
