// parsim_force_pairs2.cuh -- force + integration with TWO targets per thread (updateForces +
// applyForce, serial/parsim.cpp:292-377,:150-195); same arithmetic and the same summation order per
// target as k_force_integrate (parsim_kernels.cuh), so the results are the same bit for bit.
//
// Why: in k_force_integrate every pair interaction costs one 16-byte and one 8-byte shared-memory
// load (the source's position and mass, broadcast to the lanes of the cell).  ncu shows the
// shared-memory pipe 43-49 % busy next to an FP64 pipe 69-78 % busy (profiles/
// ncu_force10_r01_v7_summary.txt): about 4.4 shared-memory wavefronts per warp and pair -- the two
// pipes are both close enough to their limits to get in each other's way.  Here a thread owns two
// CONSECUTIVE particles (almost always of the same cell) and uses every source it loads for both:
// half the shared-memory traffic per pair, twice the independent FP64 chains per thread.
//
// Selected with PARSIM_FORCE_VARIANT=12 / 13; has NOT run on a GPU yet, so it is not the default
// and not part of the variant-agreement test.
#pragma once
#include "parsim_kernels.cuh"

namespace parsim {

// the 8 neighbour centres of mass of local cell c acting on p, in the reference's order
// (dx outer, dy inner, :301-305), shifted by +-side across the periodic border (:314-329)
__device__ __forceinline__ void neighbours_direct(const Geom &g, int c, double2 p, const double2 *__restrict__ com_xy,
                                                  const double *__restrict__ com_m, double &ax, double &ay) {
    const int lrow = c / g.nc;
    const int gx = c - lrow * g.nc;
    const int grow = g.row0 + lrow;
#pragma unroll
    for (int dxn = -1; dxn <= 1; dxn++) {
        int nx = gx + dxn;
        double sx = 0.0;
        if (nx >= g.nc) { sx = g.side; nx -= g.nc; } else if (nx < 0) { sx = -g.side; nx += g.nc; }
#pragma unroll
        for (int dyn = -1; dyn <= 1; dyn++) {
            if (dxn == 0 && dyn == 0) continue;
            const int ny = grow + dyn;
            double sy = 0.0;
            if (ny >= g.nc) sy = g.side; else if (ny < 0) sy = -g.side;
            const int q = (lrow + dyn + 1) * g.nc + nx;
            const double2 cm = __ldg(&com_xy[q]);
            const double M = __ldg(&com_m[q]);
            accumulate((sx + cm.x) - p.x, (sy + cm.y) - p.y, M, ax, ay);
        }
    }
}

// Particle::updatePositionAndVelocity + wrap + new cell + mover list for one target
// (serial/parsim.cpp:54-60,:171-190), the epilogue of k_force_integrate
__device__ __forceinline__ void integrate_store(const Geom &g, int i, int c, double2 p, double2 v, double ax, double ay,
                                                double2 *__restrict__ pos_out, float2 *__restrict__ posf_out,
                                                double2 *__restrict__ vel, int *__restrict__ newcell,
                                                unsigned char *__restrict__ flags, int *__restrict__ mover_list,
                                                int *n_movers, double2 *__restrict__ dbg_acc) {
    ax *= kG;
    ay *= kG;
    if (dbg_acc) dbg_acc[i] = make_double2(ax, ay);
    double x = __dadd_rn(p.x, __dadd_rn(__dmul_rn(v.x, kDeltaT),
                                        __dmul_rn(__dmul_rn(__dmul_rn(0.5, ax), kDeltaT), kDeltaT)));
    double y = __dadd_rn(p.y, __dadd_rn(__dmul_rn(v.y, kDeltaT),
                                        __dmul_rn(__dmul_rn(__dmul_rn(0.5, ay), kDeltaT), kDeltaT)));
    v.x = __dadd_rn(v.x, __dmul_rn(ax, kDeltaT));
    v.y = __dadd_rn(v.y, __dmul_rn(ay, kDeltaT));
    x = wrap_coord(x, g.side);
    y = wrap_coord(y, g.side);
    const double2 pn = make_double2(x, y);
    pos_out[i] = pn;
    if (posf_out) posf_out[i] = make_float2((float)x, (float)y);
    vel[i] = v;
    const int cnew = global_cell_fast(pn, g);
    newcell[i] = cnew;
    const bool mover = cnew != c + g.cell_base;
    flags[i] = mover ? (unsigned char)F_MOVER : (unsigned char)0;
    if (mover) {
        const unsigned mm = __activemask();
        const int lane = threadIdx.x & 31, leader = __ffs(mm) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(n_movers, __popc(mm));
        base = __shfl_sync(mm, base, leader);
        mover_list[base + __popc(mm & ((1u << lane) - 1u))] = i;
    }
}

// block = TB threads = 2 TB consecutive particles; TILE staged sources, CELLS cells whose shifted
// neighbourhood is staged (as in k_force_integrate)
template <int TB, int UNROLL, int TILE, int CELLS, int MINB>
__global__ void __launch_bounds__(TB, MINB)
k_force_integrate2(Geom g, const int *__restrict__ n_ptr, const double2 *__restrict__ pos_in, const double *__restrict__ m,
                   const int *__restrict__ cell, const int *__restrict__ cell_start,
                   const double2 *__restrict__ com_xy, const double *__restrict__ com_m,
                   double2 *__restrict__ pos_out, float2 *__restrict__ posf_out, double2 *__restrict__ vel,
                   int *__restrict__ newcell, unsigned char *__restrict__ flags, int *__restrict__ mover_list,
                   int *n_movers, double2 *__restrict__ dbg_acc) {
    __shared__ double2 s_pos[TILE];
    __shared__ double s_m[TILE];
    __shared__ double2 s_nxy[CELLS * 8];
    __shared__ double s_nm[CELLS * 8];
    const int n = *n_ptr;
    const int b0 = blockIdx.x * (2 * TB);
    if (b0 >= n) return;
    const int i0 = b0 + 2 * threadIdx.x, i1 = i0 + 1;
    const int last = min(b0 + 2 * TB, n) - 1;
    const int cf = cell[b0], cl = cell[last];
    const int r0 = cell_start[cf], r1 = cell_start[cl + 1];
    const int ncb = cl - cf + 1;
    const bool staged = r1 - r0 <= TILE;  // block-uniform
    const bool nb_staged = ncb <= CELLS;  // block-uniform
    const bool ok0 = i0 < n, ok1 = i1 < n;
    int c0 = 0, c1 = 0, s0 = 0, e0 = 0, s1 = 0, e1 = 0;
    double2 p0 = make_double2(0.0, 0.0), p1 = p0;
    if (ok0) { c0 = cell[i0]; p0 = pos_in[i0]; s0 = cell_start[c0]; e0 = cell_start[c0 + 1]; }
    if (ok1) { c1 = cell[i1]; p1 = pos_in[i1]; s1 = cell_start[c1]; e1 = cell_start[c1 + 1]; }
    if (staged) {
        for (int k = r0 + threadIdx.x; k < r1; k += TB) {
            s_pos[k - r0] = pos_in[k];
            s_m[k - r0] = m[k];
        }
    }
    if (nb_staged) {
        for (int idx = threadIdx.x; idx < ncb * 8; idx += TB) {
            const int cc = cf + (idx >> 3), j = idx & 7;
            const int dxn = j < 3 ? -1 : j < 5 ? 0 : 1;
            const int dyn = j < 3 ? j - 1 : j == 3 ? -1 : j == 4 ? 1 : j - 6;
            const int lrow = cc / g.nc;
            int nx = cc - lrow * g.nc + dxn;
            double sx = 0.0, sy = 0.0;
            if (nx >= g.nc) { sx = g.side; nx -= g.nc; } else if (nx < 0) { sx = -g.side; nx += g.nc; }
            const int ny = g.row0 + lrow + dyn;
            if (ny >= g.nc) sy = g.side; else if (ny < 0) sy = -g.side;
            const int q = (lrow + dyn + 1) * g.nc + nx;
            const double2 cm = __ldg(&com_xy[q]);
            s_nxy[idx] = make_double2(sx + cm.x, sy + cm.y);
            s_nm[idx] = __ldg(&com_m[q]);
        }
    }
    double2 v0 = make_double2(0.0, 0.0), v1 = v0;
    if (ok0) v0 = vel[i0];
    if (ok1) v1 = vel[i1];
    if (staged || nb_staged) __syncthreads();
    if (!ok0) return;

    double ax0 = 0.0, ay0 = 0.0, ax1 = 0.0, ay1 = 0.0;
    // the 8 neighbour centres of mass first, as in k_force_integrate
    if (nb_staged) {
        const int base0 = (c0 - cf) * 8, base1 = (c1 - cf) * 8;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const double2 cm = s_nxy[base0 + j];
            accumulate(cm.x - p0.x, cm.y - p0.y, s_nm[base0 + j], ax0, ay0);
        }
        if (ok1) {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const double2 cm = s_nxy[base1 + j];
                accumulate(cm.x - p1.x, cm.y - p1.y, s_nm[base1 + j], ax1, ay1);
            }
        }
    } else {
        neighbours_direct(g, c0, p0, com_xy, com_m, ax0, ay0);
        if (ok1) neighbours_direct(g, c1, p1, com_xy, com_m, ax1, ay1);
    }
    // then the own cell's particles in list order; one pass serves both targets when they share it
    if (ok1 && c1 == c0) {
        if (staged) {
#pragma unroll UNROLL
            for (int k = s0 - r0; k < e0 - r0; k++) {
                const double2 q = s_pos[k];
                const double mk = s_m[k];
                accumulate(q.x - p0.x, q.y - p0.y, mk, ax0, ay0);
                accumulate(q.x - p1.x, q.y - p1.y, mk, ax1, ay1);
            }
        } else {
#pragma unroll 2
            for (int k = s0; k < e0; k++) {
                const double2 q = __ldg(&pos_in[k]);
                const double mk = __ldg(&m[k]);
                accumulate(q.x - p0.x, q.y - p0.y, mk, ax0, ay0);
                accumulate(q.x - p1.x, q.y - p1.y, mk, ax1, ay1);
            }
        }
    } else {
        if (staged) {
#pragma unroll UNROLL
            for (int k = s0 - r0; k < e0 - r0; k++) {
                const double2 q = s_pos[k];
                accumulate(q.x - p0.x, q.y - p0.y, s_m[k], ax0, ay0);
            }
            if (ok1) {
#pragma unroll UNROLL
                for (int k = s1 - r0; k < e1 - r0; k++) {
                    const double2 q = s_pos[k];
                    accumulate(q.x - p1.x, q.y - p1.y, s_m[k], ax1, ay1);
                }
            }
        } else {
#pragma unroll 2
            for (int k = s0; k < e0; k++) {
                const double2 q = __ldg(&pos_in[k]);
                accumulate(q.x - p0.x, q.y - p0.y, __ldg(&m[k]), ax0, ay0);
            }
            if (ok1) {
#pragma unroll 2
                for (int k = s1; k < e1; k++) {
                    const double2 q = __ldg(&pos_in[k]);
                    accumulate(q.x - p1.x, q.y - p1.y, __ldg(&m[k]), ax1, ay1);
                }
            }
        }
    }
    integrate_store(g, i0, c0, p0, v0, ax0, ay0, pos_out, posf_out, vel, newcell, flags, mover_list, n_movers, dbg_acc);
    if (ok1)
        integrate_store(g, i1, c1, p1, v1, ax1, ay1, pos_out, posf_out, vel, newcell, flags, mover_list, n_movers, dbg_acc);
}

}  // namespace parsim
