// parsim_force_tiles.cuh -- persistent, software-pipelined instantiation of the force + integration
// kernel (updateForces + applyForce, serial/parsim.cpp:292-377,:150-195); same arithmetic, same
// summation order and same outputs as k_force_integrate (parsim_kernels.cuh), bit for bit.
//
// Why: ncu (profiles/ncu_force10_r01_v7_raw.csv) shows k_force_integrate's FP64 pipe 69-78 % busy
// with, per scheduler, about 2 of 7 resident warps parked at the staging __syncthreads and 1.4-3
// more waiting for the global loads that precede it: every block pays its own load -> barrier ->
// pair loop -> store sequence.  Here a block stays resident and walks over many 128-target tiles;
// a PRODUCER warp runs up to NST - 1 tiles ahead of the four CONSUMER warps and stages everything a
// tile needs in one of NST shared-memory buffers:
//   * the contiguous source range of the tile (positions, masses), the tile's cell indices,
//     velocities and the window of cell_start it needs -- five 1-D bulk copies of the TMA engine
//     (cp.async.bulk global -> shared, completion counted in bytes on an mbarrier);
//   * the 8 neighbour centres of mass of every cell of the tile, already shifted across the
//     periodic border (the 3 x 3 neighbourhood of north_star (c)), written by the producer lanes.
// Consumers wait on the buffer's "full" mbarrier, run the pair loop out of shared memory, hand the
// buffer back on its "empty" mbarrier and only then do the integration and the global stores, so
// neither a block-wide barrier nor a global load sits in front of the FP64-bound loop.
//
// Selected with PARSIM_FORCE_VARIANT=5 (pair loop unrolled 4 times, two 768-source buffers, 4 blocks
// per SM at 96 registers), 6 (unrolled 8 times), 7 (two 448-source buffers, 6 blocks per SM at 64
// registers), 8 (three 448-source buffers, 4 blocks per SM) or 9 (as 5 with one mbarrier arrival per
// warp); 8 and 9 have not run on a GPU yet.  Measured on B200
// (config 3, profiles/bench_r01_v11_fv*.json): results identical to the default (all GPU parity
// tests, 29 214 105 collisions at 1e8 particles), but 33.1 / 32.7 /
// 32.8 ms per run for variants 5 / 6 / 7 against 28.8 ms of k_force_integrate.  The slowdown is the
// same 13-16 % in EVERY timestep (profiles/force_steps_r01_v11.txt: 11.7 against 10.3 ms in the
// full first step, 1.8 against 1.5 ms in the thinnest), and the three settings agree although
// their occupancy (16 / 16 / 24 consumer warps per SM against the default's 32) and unrolling
// differ; the unrolled pair loops are the same SASS (74 against 72 instructions per 4 pairs).  So
// neither the producer's latency nor the loop body explains it, and parking the waiting warps
// (try_wait with a suspend-time hint, nanosleep back-off in the producer) instead of spinning on
// the mbarriers changed nothing (force_steps_r01_v12.txt); an ncu capture of this kernel is the
// next step.  NOT the default.
#pragma once
#include "parsim_kernels.cuh"

namespace parsim {

constexpr int kFtTargets = 128;               // targets per tile = consumer threads per block
constexpr int kFtThreads = kFtTargets + 32;   // + the producer warp
constexpr int kFtCells = 16;                  // cells per tile whose neighbourhood is staged
constexpr int kFtCs = 144;                    // staged cell_start entries (16-byte aligned window)

// The bulk copies move whole 16-byte units from 16-byte aligned addresses, so the windows of the
// 8-byte and 4-byte arrays start at an even / multiple-of-4 element and may run a few elements
// past the part that is used; rank_create pads m, cell and cell_start accordingly (kFtPad).
constexpr int kFtPad = 8;

template <int TILE>  // staged sources per tile
struct __align__(16) FtBuf {
    double2 pos[TILE];           // sources r0 .. r1
    double2 vel[kFtTargets];     // velocities of the tile's targets
    double2 nxy[kFtCells * 8];   // shifted neighbour centres of mass, 8 per cell of the tile
    double m[TILE + 2];          // masses from m_base = r0 rounded down to even
    double nm[kFtCells * 8];
    int cell[kFtTargets];        // cells of the tile's targets
    int cs[kFtCs];               // cell_start from cs_base = first cell rounded down to a multiple of 4
    int r0, r1, m_base, cs_base, cf, ncb, flags, pad;
};
static_assert(sizeof(FtBuf<768>) % 16 == 0 && sizeof(FtBuf<448>) % 16 == 0,
              "FtBuf must keep its members 16-byte aligned in an array");
enum { FT_SRC = 1, FT_NB = 2, FT_CS = 4 };  // FtBuf::flags: what was staged for this tile

template <int TILE, int NST>  // NST buffers: the producer runs up to NST - 1 tiles ahead
inline size_t force_tiles_smem() { return NST * sizeof(FtBuf<TILE>) + 2 * NST * sizeof(unsigned long long); }

__device__ __forceinline__ unsigned ft_smem(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ft_smem(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ft_smem(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ft_smem(bar)), "r"(bytes) : "memory");
}
// returns once the phase of the given parity has completed.  The suspend-time hint lets the
// hardware park the warp until the phase completes instead of returning early to poll again.
__device__ __forceinline__ bool mbar_try_wait(unsigned a, unsigned parity) {
    unsigned done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(a), "r"(parity), "r"(0x989680u) : "memory");
    return done != 0;
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    const unsigned a = ft_smem(bar);
    while (!mbar_try_wait(a, parity)) {}
}
// the producer waits a whole tile's worth of pair work: back off between polls so that its
// mbarrier traffic does not compete with the consumers' shared-memory loads
__device__ __forceinline__ void mbar_wait_backoff(unsigned long long *bar, unsigned parity) {
    const unsigned a = ft_smem(bar);
    while (!mbar_try_wait(a, parity)) __nanosleep(200);
}
// 1-D bulk copy global -> shared (TMA engine); bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(ft_smem(dst)), "l"((unsigned long long)__cvta_generic_to_global(src)), "r"(bytes),
                   "r"(ft_smem(bar)) : "memory");
}

// ELECT: one arrival per WARP on the mbarriers (after __syncwarp, which orders the lanes' shared-memory
// accesses before the elected lane's release) instead of one per thread: 5 instead of 160 mbarrier
// operations per tile.
template <int UNROLL, int TILE, int MINB, int NST, bool ELECT = false>
__global__ void __launch_bounds__(kFtThreads, MINB)
k_force_tiles(Geom g, int n, const double2 *__restrict__ pos_in, const double *__restrict__ m,
              const int *__restrict__ cell, const int *__restrict__ cell_start,
              const double2 *__restrict__ com_xy, const double *__restrict__ com_m,
              double2 *__restrict__ pos_out, float2 *__restrict__ posf_out, double2 *vel,
              int *__restrict__ newcell, unsigned char *__restrict__ flags, int *__restrict__ mover_list,
              int *n_movers, double2 *__restrict__ dbg_acc) {
    extern __shared__ __align__(16) unsigned char ft_raw[];
    typedef FtBuf<TILE> Buf;
    Buf *buf = reinterpret_cast<Buf *>(ft_raw);
    unsigned long long *full = reinterpret_cast<unsigned long long *>(ft_raw + NST * sizeof(Buf));
    unsigned long long *empty = full + NST;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = (n + kFtTargets - 1) / kFtTargets;
    if (threadIdx.x == 0) {
        for (int b = 0; b < NST; b++) {
            mbar_init(&full[b], ELECT ? 1 : 32);  // the producer lanes (+ the bytes of the bulk copies)
            mbar_init(&empty[b], ELECT ? kFtTargets / 32 : kFtTargets);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp == kFtTargets / 32) {
        // ---------------- producer warp: one to two tiles ahead of the consumers ----------------
        int j = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, j++) {
            const int b = j % NST, use = j / NST;
            Buf &B = buf[b];
            if (use > 0) mbar_wait_backoff(&empty[b], (unsigned)((use - 1) & 1));  // consumers are done with it
            const int b0 = tile * kFtTargets;
            const int cnt = min(kFtTargets, n - b0);
            const int cf = __ldg(&cell[b0]), cl = __ldg(&cell[b0 + cnt - 1]);
            const int r0 = __ldg(&cell_start[cf]), r1 = __ldg(&cell_start[cl + 1]);
            const int ncb = cl - cf + 1;
            const int m_base = r0 & ~1, m_cnt = ((r1 + 1) & ~1) - m_base;
            const int cs_base = cf & ~3, cs_cnt = ((cl + 2 + 3) & ~3) - cs_base;
            const int cell_cnt = (cnt + 3) & ~3;
            const bool src_staged = r1 - r0 <= TILE;
            const bool nb_staged = ncb <= kFtCells;
            const bool cs_staged = cs_cnt <= kFtCs;
            if (nb_staged) {
                // same arithmetic as k_force_integrate: neighbour j of cell cc in the reference's order
                // (dx outer, dy inner, :301-305), shifted by +-side across the border (:314-329)
                for (int idx = lane; idx < ncb * 8; idx += 32) {
                    const int cc = cf + (idx >> 3), jn = idx & 7;
                    const int dxn = jn < 3 ? -1 : jn < 5 ? 0 : 1;
                    const int dyn = jn < 3 ? jn - 1 : jn == 3 ? -1 : jn == 4 ? 1 : jn - 6;
                    const int lrow = cc / g.nc;
                    int nx = cc - lrow * g.nc + dxn;
                    double sx = 0.0, sy = 0.0;
                    if (nx >= g.nc) { sx = g.side; nx -= g.nc; } else if (nx < 0) { sx = -g.side; nx += g.nc; }
                    const int ny = g.row0 + lrow + dyn;
                    if (ny >= g.nc) sy = g.side; else if (ny < 0) sy = -g.side;
                    const int q = (lrow + dyn + 1) * g.nc + nx;
                    const double2 cm = __ldg(&com_xy[q]);
                    B.nxy[idx] = make_double2(sx + cm.x, sy + cm.y);
                    B.nm[idx] = __ldg(&com_m[q]);
                }
            }
            if (ELECT) __syncwarp();  // the lanes' nxy / nm stores come before lane 0's release
            if (lane == 0) {
                B.r0 = r0; B.r1 = r1; B.m_base = m_base; B.cs_base = cs_base; B.cf = cf; B.ncb = ncb;
                B.flags = (src_staged ? FT_SRC : 0) | (nb_staged ? FT_NB : 0) | (cs_staged ? FT_CS : 0);
                const unsigned bytes = 4u * cell_cnt + 16u * cnt + (src_staged ? 16u * (r1 - r0) + 8u * m_cnt : 0u) +
                                       (cs_staged ? 4u * cs_cnt : 0u);
                mbar_arrive_expect_tx(&full[b], bytes);
                bulk_g2s(B.cell, cell + b0, 4u * cell_cnt, &full[b]);
                bulk_g2s(B.vel, vel + b0, 16u * cnt, &full[b]);
                if (src_staged) {
                    bulk_g2s(B.pos, pos_in + r0, 16u * (r1 - r0), &full[b]);
                    bulk_g2s(B.m, m + m_base, 8u * m_cnt, &full[b]);
                }
                if (cs_staged) bulk_g2s(B.cs, cell_start + cs_base, 4u * cs_cnt, &full[b]);
            } else if (!ELECT) {
                mbar_arrive(&full[b]);  // releases this lane's nxy / nm stores
            }
        }
        return;
    }

    // -------------------------------- consumer warps ------------------------------------------
    int j = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, j++) {
        const int b = j % NST, use = j / NST;
        const Buf &B = buf[b];
        mbar_wait(&full[b], (unsigned)(use & 1));
        const int i = tile * kFtTargets + threadIdx.x;
        const bool valid = i < n;
        int c = 0;
        double2 p = make_double2(0.0, 0.0), v = make_double2(0.0, 0.0);
        double ax = 0.0, ay = 0.0;
        if (valid) {
            const int fl = B.flags, r0 = B.r0;
            c = B.cell[threadIdx.x];
            v = B.vel[threadIdx.x];
            int s, e;
            if (fl & FT_CS) { s = B.cs[c - B.cs_base]; e = B.cs[c - B.cs_base + 1]; }
            else { s = __ldg(&cell_start[c]); e = __ldg(&cell_start[c + 1]); }
            p = (fl & FT_SRC) ? B.pos[i - r0] : pos_in[i];
            if (fl & FT_NB) {
                const int base = (c - B.cf) * 8;
#pragma unroll
                for (int jn = 0; jn < 8; jn++) {
                    const double2 cm = B.nxy[base + jn];
                    accumulate(cm.x - p.x, cm.y - p.y, B.nm[base + jn], ax, ay);
                }
            } else {  // the tile touches too many cells (sparse grid): through L1, as in k_force_integrate
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
            if (fl & FT_SRC) {
                const double2 *sp = B.pos - r0;
                const double *sm = B.m - B.m_base;
#pragma unroll UNROLL
                for (int k = s; k < e; k++) {
                    const double2 q = sp[k];
                    accumulate(q.x - p.x, q.y - p.y, sm[k], ax, ay);
                }
            } else {  // very full cells: the range does not fit the buffer
#pragma unroll 4
                for (int k = s; k < e; k++) {
                    const double2 q = __ldg(&pos_in[k]);
                    accumulate(q.x - p.x, q.y - p.y, __ldg(&m[k]), ax, ay);
                }
            }
        }
        // every shared-memory read of this tile is behind us
        if (ELECT) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[b]);
        } else {
            mbar_arrive(&empty[b]);
        }
        if (!valid) continue;
        ax *= kG;
        ay *= kG;
        if (dbg_acc) dbg_acc[i] = make_double2(ax, ay);

        // Particle::updatePositionAndVelocity, serial/parsim.cpp:54-60, same rounding sequence
        double x = __dadd_rn(p.x, __dadd_rn(__dmul_rn(v.x, kDeltaT),
                                            __dmul_rn(__dmul_rn(__dmul_rn(0.5, ax), kDeltaT), kDeltaT)));
        double y = __dadd_rn(p.y, __dadd_rn(__dmul_rn(v.y, kDeltaT),
                                            __dmul_rn(__dmul_rn(__dmul_rn(0.5, ay), kDeltaT), kDeltaT)));
        v.x = __dadd_rn(v.x, __dmul_rn(ax, kDeltaT));
        v.y = __dadd_rn(v.y, __dmul_rn(ay, kDeltaT));
        x = wrap_coord(x, g.side);  // :171-172
        y = wrap_coord(y, g.side);
        const double2 pn = make_double2(x, y);
        pos_out[i] = pn;
        if (posf_out) posf_out[i] = make_float2((float)x, (float)y);
        vel[i] = v;
        const int cnew = global_cell_fast(pn, g);  // :174-177
        newcell[i] = cnew;
        const bool mover = cnew != c + g.cell_base;  // :183-190
        flags[i] = mover ? (unsigned char)F_MOVER : (unsigned char)0;
        if (mover) {
            const unsigned mm = __activemask();
            const int leader = __ffs(mm) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(n_movers, __popc(mm));
            base = __shfl_sync(mm, base, leader);
            mover_list[base + __popc(mm & ((1u << lane) - 1u))] = i;
        }
    }
}

}  // namespace parsim
