// parsim_kernels.cuh -- hand-written FP64 CUDA kernels (sm_100a) for the per-timestep loop of
// `parsim` (reference: serial/parsim.cpp:422-448).  No tensor cores: nothing here is a dense
// contraction.  Layout, roofline and algorithmic bytes of every kernel are in DESIGN.md.
//
// Device data model (one row strip of the grid per GPU):
//   alive particles only, sorted by (cell, id):  pos (double2 x,y), vel (double2 vx,vy),
//   m (double), id (u32), cell (local cell, int);  cell_start[C+1] is the CSR over cells.
//   Dead particles are compacted out every step and leave a tombstone (cell, id) in a second
//   CSR, because the reference keeps them in its per-cell lists where they decide which movers
//   the list update leaves behind (serial/parsim.cpp:242-254).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace parsim {

// serial/parsim.cpp:13-16
constexpr double kG = 6.67408e-11;
constexpr double kEpsilon = 0.005;
constexpr double kDeltaT = 0.1;

constexpr int kThreads = 256;

// per-particle flags written by the list-update kernel
enum : unsigned char {
    F_MOVER = 1,    // position cell changed in this step (serial/parsim.cpp:183)
    F_REMOVED = 2,  // list update moved it to the new cell's list (:250-251); mover && !removed = stray
    F_EMIGRANT = 4, // removed mover whose new cell belongs to another strip (sent before collisions)
    F_GUEST = 8,    // arrived from another strip this step (lives past n_alive until the rebin)
};

struct Geom {
    double side;   // side_length
    double cellw;  // side_length / grid_size, the divisor of serial/parsim.cpp:160-161,:268-269
    double inv_cellw;  // 1 / cellw (fast path of cell_coord_fast only)
    int nc;        // grid_size (cells per side)
    int row0;      // first global grid row of this strip
    int rows;      // rows in this strip
    int ncells;    // rows * nc
    int cell_base; // row0 * nc : global cell index = local + cell_base
    int no_strays; // 1 = list update of the OpenMP variant (openMP/parsim-omp.cpp:248-278): every
                   // leaver is moved; 0 = the serial erase-then-increment behaviour (:234-259)
};

struct Counters {  // one per context, device memory, copied to the host once per step
    unsigned long long collisions;
    unsigned long long force_pairs;
    unsigned long long collision_pairs;
    unsigned long long movers;
    unsigned long long strays;
    int deaths_step;     // particles that died in this step's collision phase
    int grave_count;     // entries in the graveyard (all dead so far, this strip)
    int n_emigrants_up;  // removed movers sent to the strip above / below (pre-collision)
    int n_emigrants_dn;
    int n_late_up;       // surviving strays + tombstones sent after the collision phase
    int n_late_dn;
    int n_next;          // alive residents of the strip: set at upload / reset, then by every step's rebin
                         // scan.  Kernels read the particle count from here, so a step can be enqueued
                         // before the host knows how many particles died in the previous one
    int error;           // sticky device-side error code (capacity overflow, long jump, ...)
    int n_staged;        // upload cursor for multi-rank ingest
    int n_movers;        // particles that changed cell in this step (length of the mover list)
};

// Arrival record: what the collision search needs to know about a mover that the list update
// appended to a cell's list, written next to the linked bag so that a cell's first kArrSlots
// arrivals cost one load instead of a pointer chase.
constexpr int kArrSlots = 4;
struct __align__(16) ArrRec {
    double2 pos;             // post-move position
    unsigned long long key;  // list key (see list_key)
    int idx;                 // entry index
    int pad;
};

enum DeviceError { DE_NONE = 0, DE_CAPACITY = 1, DE_LONG_JUMP = 2, DE_MIGRANT_OVERFLOW = 3 };

// ---------------------------------------------------------------------------------------
// cell of a position: (int)(x / (side/ncside)) with a true IEEE division by the quotient,
// serial/parsim.cpp:160-163,:174-177,:268-272.  The reference leaves an out-of-range cell
// to a [PANIC] print / UB (:276-280); here it is clamped so memory stays safe.
__device__ __forceinline__ int cell_coord(double v, const Geom &g) {
    int c = __double2int_rz(__ddiv_rn(v, g.cellw));
    return min(max(c, 0), g.nc - 1);
}
__device__ __forceinline__ int global_cell(double2 p, const Geom &g) {
    return cell_coord(p.y, g) * g.nc + cell_coord(p.x, g);
}

// stats helper: one global atomic per block instead of one per thread
__device__ __forceinline__ void block_add_u64(unsigned long long v, unsigned long long *dst) {
    __shared__ unsigned long long acc;
    if (threadIdx.x == 0) acc = 0;
    __syncthreads();
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&acc, v);
    __syncthreads();
    if (threadIdx.x == 0 && acc) atomicAdd(dst, acc);
}

// ---------------------------------------------------------------------------------------
// The uploaded (unordered) state: one 64-byte record per particle slot, so that the step-0 sort
// moves a particle with ONE random 64-byte access instead of four 16-byte ones (DRAM fetches
// 64 bytes at a time).  The sort never modifies the records, so parsim_reset costs nothing.
struct __align__(16) InitRec {
    double2 pos;
    double2 vel;
    double m;
    unsigned id;
    unsigned pad0;
    double2 pad1;
};
static_assert(sizeof(InitRec) == 64, "InitRec must be one 64-byte line");

__device__ __forceinline__ void store_rec(InitRec *rec, int slot, double2 p, double2 v, double m, unsigned id) {
    double2 *q = reinterpret_cast<double2 *>(rec + slot);
    q[0] = p;
    q[1] = v;
    q[2] = make_double2(m, __hiloint2double(0, (int)id));  // id in the low word of the 4th double
}

// Upload on a single strip: interleave host SoA (x,y,vx,vy,m) into records, slot = id.
// (Several strips: k_ingest_strip below keeps the strip's particles, in id order.)
__global__ void k_ingest(Geom g, long long id_base, const unsigned *__restrict__ ids, int count,
                         const double *__restrict__ x,
                         const double *__restrict__ y, const double *__restrict__ vx,
                         const double *__restrict__ vy, const double *__restrict__ m,
                         int capacity, InitRec *rec, Counters *ctr) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const unsigned pid = ids ? ids[t] : (unsigned)(id_base + t);
    if (pid >= (unsigned)capacity) { ctr->error = DE_CAPACITY; return; }
    // vx == nullptr: the velocities arrive later (k_rec_vel), see rank_upload_chunk
    store_rec(rec, (int)pid, make_double2(x[t], y[t]), vx ? make_double2(vx[t], vy[t]) : make_double2(0.0, 0.0), m[t], pid);
}

// Late velocities of a single-strip upload (slot = id): into the records ...
__global__ void k_rec_vel(int n, const double *__restrict__ vx, const double *__restrict__ vy, InitRec *rec) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) reinterpret_cast<double2 *>(rec + i)[1] = make_double2(vx[i], vy[i]);
}
// ... and into the (cell, id)-ordered velocity array: where[slot] = position of the particle in the
// ordered arrays (built while the velocities are still on the link), then every arriving piece of
// (vx, vy) is scattered through it
__global__ void k_slot_positions(int n, const unsigned *__restrict__ id, int *__restrict__ where) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) where[id[p]] = p;
}
__global__ void k_vel_scatter(int first, int count, const int *__restrict__ where, const double *__restrict__ vx,
                              const double *__restrict__ vy, double2 *__restrict__ vel) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const int s = first + t;
    vel[where[s]] = make_double2(vx[s], vy[s]);
}

// ---------------------------------------------------------------------------------------
// Device-side init_particles for seeds >= 0 (serial/parsim.cpp:220-232 with the uniform
// generator :26-32).  xorshift32 is linear over GF(2): state_{k+1} = M state_k with a 32x32 bit
// matrix M, so any thread can jump to its own offset of the ONE sequential stream.  Thread t
// generates particles t, t+T, t+2T, ...: it starts at M^(5t) s0 (binary method over the
// precomputed powers M^(2^j)) and advances by M^(5T - 5) between particles.  Matrices are stored
// by columns (col[b] = M e_b), so M v is the xor of the columns selected by the bits of v.
struct XsMatrix { unsigned col[32]; };
struct XsJumpTable {
    XsMatrix pow2[34];  // M^(2^j)
    XsMatrix stride;    // M^(5T - 5): from the state after a particle's 5 draws to the next one's
};

__device__ __forceinline__ unsigned xs_apply(const XsMatrix &A, unsigned v) {
    unsigned r = 0;
#pragma unroll
    for (int b = 0; b < 32; b++) r ^= ((v >> b) & 1u) ? A.col[b] : 0u;
    return r;
}

__device__ __forceinline__ double xs_draw(unsigned &s) {
    unsigned before = s;
    s ^= s << 13;
    s ^= s >> 17;
    s ^= s << 5;
    int sum = (int)(before + s);
    return __dadd_rn(0.5, __dmul_rn(0.2328306e-09, (double)sum));  // no FMA: reference rounding
}

// (single strip: slot = id; several strips use k_init_strip below)
__global__ void k_init_particles(Geom g, const XsJumpTable *__restrict__ jt, unsigned s0, long long n_part,
                                 InitRec *rec) {
    const long long T = (long long)gridDim.x * blockDim.x;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_part) return;
    unsigned s = s0;
    unsigned long long jump = 5ull * (unsigned long long)t;
    for (int j = 0; jump; j++, jump >>= 1)
        if (jump & 1ull) s = xs_apply(jt->pow2[j], s);
    const double ncd = (double)g.nc;
    const double cc = (double)((long long)g.nc * g.nc);
    const double nd = (double)n_part;
    for (long long i = t; i < n_part; i += T) {
        double x = __dmul_rn(xs_draw(s), g.side);
        double y = __dmul_rn(xs_draw(s), g.side);
        double vx = __ddiv_rn(__ddiv_rn(__dmul_rn(__dadd_rn(xs_draw(s), -0.5), g.side), ncd), 5.0);
        double vy = __ddiv_rn(__ddiv_rn(__dmul_rn(__dadd_rn(xs_draw(s), -0.5), g.side), ncd), 5.0);
        double m = __dmul_rn(__ddiv_rn(__ddiv_rn(__dmul_rn(__dmul_rn(xs_draw(s), 0.01), cc), nd), kG),
                             0.005 * 0.005);
        s = xs_apply(jt->stride, s);
        store_rec(rec, (int)i, make_double2(x, y), make_double2(vx, vy), m, (unsigned)i);
    }
}

// FP64 FMA-chain microbenchmark: the measured FP64 peak that the force / collision kernels are
// compared against (no such number is in MEASURED_PEAKS.json).
__global__ void k_fp64_peak(int iters, double *out) {
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5,
           a6 = a0 + 6, a7 = a0 + 7;
    const double b = 0.999999, c = 1e-9;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    double r = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (r == 123.456) out[0] = r;
}

// ---------------------------------------------------------------------------------------
// Device-wide exclusive scan of int counts (three small kernels; the data is 4 B per CELL,
// not per particle).  out has N+1 entries, out[N] = total.
constexpr int kScanItems = 8;
constexpr int kScanTile = kThreads * kScanItems;

__device__ __forceinline__ int block_exclusive_scan(int v, int *warp_sums, int &block_total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < (kThreads / 32) ? warp_sums[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        if (lane < (kThreads / 32)) warp_sums[lane] = winc - w;  // exclusive warp offsets
        if (lane == 31) warp_sums[32] = winc;                    // block total
    }
    __syncthreads();
    block_total = warp_sums[32];
    int r = warp_sums[warp] + inc - v;
    __syncthreads();
    return r;
}

__global__ void k_scan_tile_sums(const int *__restrict__ in, const int *__restrict__ in2, int N,
                                 int *tile_sums) {
    __shared__ int ws[33];
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int s = 0;
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        if (k < N) s += in[k] + (in2 ? in2[k] : 0);
    }
    int tot;
    block_exclusive_scan(s, ws, tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

__global__ void k_scan_top(int *tile_sums, int ntiles, int *total_out, int *total_out2) {
    __shared__ int ws[33];
    int carry = 0;
    for (int base = 0; base < ntiles; base += kThreads) {
        int k = base + threadIdx.x;
        int v = k < ntiles ? tile_sums[k] : 0;
        int tot;
        int ex = block_exclusive_scan(v, ws, tot);
        if (k < ntiles) tile_sums[k] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) {
        *total_out = carry;
        if (total_out2) *total_out2 = carry;
    }
}

// in2 (optional) is added element-wise to in: used for tombstones (old + new counts)
__global__ void k_scan_apply(const int *__restrict__ in, const int *__restrict__ in2, int N,
                             const int *__restrict__ tile_sums, int *__restrict__ out) {
    __shared__ int ws[33];
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int v[kScanItems];
    int s = 0;
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        v[j] = k < N ? in[k] + (in2 ? in2[k] : 0) : 0;
        s += v[j];
    }
    int tot;
    int ex = block_exclusive_scan(s, ws, tot) + tile_sums[blockIdx.x];
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        if (k < N) out[k] = ex;
        ex += v[j];
    }
}

// ---------------------------------------------------------------------------------------
// Row strips (several GPUs): every strip sees the whole particle stream and keeps the particles
// whose row it owns -- a STABLE compaction (count per block, scan, write), so the kept records
// stay in id order and the step-0 sort needs no id passes.
//   PHASE 0: blk_cnt[block] = particles kept by the block
//   PHASE 1: records written at base + blk_off[block] + (rank inside the block)
constexpr int kInitPer = 32;  // particles per thread in k_init_strip (block = kThreads * kInitPer ids)

template <int PHASE>
__global__ void __launch_bounds__(kThreads)
k_ingest_strip(Geom g, long long id_base, const unsigned *__restrict__ ids, int count,
               const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ vx,
               const double *__restrict__ vy, const double *__restrict__ m, int capacity,
               int *__restrict__ blk, InitRec *rec, Counters *ctr, int *__restrict__ up_slot = nullptr) {
    // up_slot (PHASE 1, optional): the slot every uploaded particle went to, -1 if another strip owns it --
    // velocities that arrive later find their records through it (vx == nullptr: they do, rank_upload_strip)
    __shared__ int ws[33];
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    bool keep = false;
    double py = 0.0;
    if (t < count) {
        py = y[t];
        const int row = cell_coord(py, g);
        keep = row >= g.row0 && row < g.row0 + g.rows;
    }
    int tot;
    const int rank = block_exclusive_scan(keep ? 1 : 0, ws, tot);
    if (PHASE == 0) {
        if (threadIdx.x == 0) blk[blockIdx.x] = tot;
        return;
    }
    if (!keep) {
        if (up_slot && t < count) up_slot[t] = -1;
        return;
    }
    const int slot = ctr->n_staged + blk[blockIdx.x] + rank;
    if (slot >= capacity) {
        ctr->error = DE_CAPACITY;
        if (up_slot) up_slot[t] = -1;
        return;
    }
    if (up_slot) up_slot[t] = slot;
    const unsigned pid = ids ? ids[t] : (unsigned)(id_base + t);
    store_rec(rec, slot, make_double2(x[t], py), vx ? make_double2(vx[t], vy[t]) : make_double2(0.0, 0.0), m[t], pid);
}

// Late velocities of a strip upload (upload order, up_slot from k_ingest_strip): into the ordered
// velocity array through where[slot] = ordered position (k_sort_gather), and into the records
__global__ void k_vel_scatter_strip(int first, int count, const int *__restrict__ up_slot, const int *__restrict__ where,
                                    const double *__restrict__ vx, const double *__restrict__ vy, double2 *__restrict__ vel) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const int s = up_slot[first + t];
    if (s >= 0) vel[where[s]] = make_double2(vx[first + t], vy[first + t]);
}
__global__ void k_rec_vel_strip(int count, const int *__restrict__ up_slot, const double *__restrict__ vx,
                                const double *__restrict__ vy, InitRec *rec) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const int s = up_slot[t];
    if (s >= 0) reinterpret_cast<double2 *>(rec + s)[1] = make_double2(vx[t], vy[t]);
}

__global__ void k_advance_staged(Counters *ctr, const int *total) { ctr->n_staged += *total; }

__device__ __forceinline__ void xs_step(unsigned &s) {
    s ^= s << 13;
    s ^= s >> 17;
    s ^= s << 5;
}

// jt->stride must be M^(5 * kThreads - 5) here: a thread's particles are kThreads ids apart.
template <int PHASE>
__global__ void __launch_bounds__(kThreads)
k_init_strip(Geom g, const XsJumpTable *__restrict__ jt, unsigned s0, long long n_part, int capacity,
             int *__restrict__ blk, InitRec *rec, Counters *ctr) {
    __shared__ int ws[33];
    const long long first = (long long)blockIdx.x * (kThreads * kInitPer) + threadIdx.x;
    unsigned s = s0;
    unsigned long long jump = 5ull * (unsigned long long)first;
    for (int j = 0; jump; j++, jump >>= 1)
        if (jump & 1ull) s = xs_apply(jt->pow2[j], s);
    const double ncd = (double)g.nc;
    const double cc = (double)((long long)g.nc * g.nc);
    const double nd = (double)n_part;
    int running = PHASE == 1 ? ctr->n_staged + blk[blockIdx.x] : 0;
    int kept = 0;
    for (int r = 0; r < kInitPer; r++) {
        const long long i = first + (long long)r * kThreads;
        const double x = __dmul_rn(xs_draw(s), g.side);
        const double y = __dmul_rn(xs_draw(s), g.side);
        const int row = cell_coord(y, g);
        const bool keep = i < n_part && row >= g.row0 && row < g.row0 + g.rows;
        if (PHASE == 0) {
            xs_step(s); xs_step(s); xs_step(s);
            kept += keep ? 1 : 0;
        } else {
            const double vx = __ddiv_rn(__ddiv_rn(__dmul_rn(__dadd_rn(xs_draw(s), -0.5), g.side), ncd), 5.0);
            const double vy = __ddiv_rn(__ddiv_rn(__dmul_rn(__dadd_rn(xs_draw(s), -0.5), g.side), ncd), 5.0);
            const double m = __dmul_rn(__ddiv_rn(__ddiv_rn(__dmul_rn(__dmul_rn(xs_draw(s), 0.01), cc), nd), kG),
                                       0.005 * 0.005);
            int tot;
            const int rank = block_exclusive_scan(keep ? 1 : 0, ws, tot);
            if (keep) {
                const int slot = running + rank;
                if (slot >= capacity) ctr->error = DE_CAPACITY;
                else store_rec(rec, slot, make_double2(x, y), make_double2(vx, vy), m, (unsigned)i);
            }
            running += tot;
        }
        s = xs_apply(jt->stride, s);
    }
    if (PHASE == 0) {
        int tot;
        block_exclusive_scan(kept, ws, tot);
        if (threadIdx.x == 0) blk[blockIdx.x] = tot;
    }
}

// ---------------------------------------------------------------------------------------
// Centre of mass per cell (north_star (b); Cell::addParticle, serial/parsim.cpp:93-106 computes
// the same weighted mean as a running average).  LANES lanes per cell, each lane sums a fixed
// strided subsequence, then a fixed xor-shuffle tree: the order of additions depends only on
// the cell's (id-sorted) contents, never on scheduling or on the number of GPUs.
// com_xy / com_m carry one ghost row above and below the strip: index = (local_row+1)*nc + gx.
constexpr int kComCells = 2;  // consecutive cells per lane group
// rounds of every cell whose loads are issued before any is used (fewer for narrow lane groups:
// they are chosen for sparse cells, where the later rounds are mostly empty)
#ifndef PS_COM_AHEAD_NARROW
#define PS_COM_AHEAD_NARROW 2
#endif

#ifndef PS_COM_MINB
#define PS_COM_MINB 4
#endif
template <int LANES>
__global__ void __launch_bounds__(kThreads, LANES <= 8 ? 8 : PS_COM_MINB) k_com(Geom g, const double2 *__restrict__ pos, const double *__restrict__ m,
                      const int *__restrict__ cell_start, double2 *__restrict__ com_xy,
                      double *__restrict__ com_m, unsigned long long *force_pairs) {
    constexpr int kComAhead = LANES <= 8 ? PS_COM_AHEAD_NARROW : 4;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int c0 = (t / LANES) * kComCells, lane = t % LANES;
    // Each lane group takes kComCells consecutive cells and has the first kComAhead rounds of all of
    // them in flight at once (memory-level parallelism: a cell is only ~2.4 KB).  Out-of-range
    // slots add exact zeros, so the order of additions is that of the plain strided loop.
    int b[kComCells + 1];
#pragma unroll
    for (int q = 0; q <= kComCells; q++) b[q] = cell_start[min(c0 + q, g.ncells)];
    double mk[kComCells][kComAhead];
    double2 pk[kComCells][kComAhead];
#pragma unroll
    for (int q = 0; q < kComCells; q++) {
#pragma unroll
        for (int j = 0; j < kComAhead; j++) {
            const int k = b[q] + lane + j * LANES;
            const bool in = k < b[q + 1];
            mk[q][j] = in ? m[k] : 0.0;
            pk[q][j] = in ? pos[k] : make_double2(0.0, 0.0);
        }
    }
    unsigned long long pairs = 0;
    double Ms[kComCells], Mxs[kComCells], Mys[kComCells];
#pragma unroll
    for (int q = 0; q < kComCells; q++) {
        const int s = b[q], e = b[q + 1];
        double M = 0, Mx = 0, My = 0;
#pragma unroll
        for (int j = 0; j < kComAhead; j++) {
            M += mk[q][j];
            Mx = fma(mk[q][j], pk[q][j].x, Mx);
            My = fma(mk[q][j], pk[q][j].y, My);
        }
        for (int k = s + lane + kComAhead * LANES; k < e; k += LANES) {
            double mv = m[k];
            double2 p = pos[k];
            M += mv;
            Mx = fma(mv, p.x, Mx);
            My = fma(mv, p.y, My);
        }
#pragma unroll
        for (int o = LANES / 2; o > 0; o >>= 1) {
            M += __shfl_xor_sync(0xffffffffu, M, o);
            Mx += __shfl_xor_sync(0xffffffffu, Mx, o);
            My += __shfl_xor_sync(0xffffffffu, My, o);
        }
        Ms[q] = M; Mxs[q] = Mx; Mys[q] = My;  // every lane of the group holds the totals
    }
    // The two divisions per cell are the long pole of this kernel (an FP64 division is ~40
    // instructions and the whole warp steps through it): lanes 0..3 of the group do the four
    // divisions of its two cells in ONE pass -- lane 2q + j divides component j of cell q.
    static_assert(kComCells == 2 && LANES >= 4, "division layout");
    {
        const int q = (lane >> 1) & 1;
        const double num = (lane & 1) ? (q ? Mys[1] : Mys[0]) : (q ? Mxs[1] : Mxs[0]);
        const double den = q ? Ms[1] : Ms[0];
        double val = 0.0;
        if (lane < 4 && den > 0) val = num / den;
        const double other = __shfl_down_sync(0xffffffffu, val, 1);
        const int c = c0 + q;
        if (lane < 4 && (lane & 1) == 0 && c < g.ncells) {
            com_xy[c + g.nc] = make_double2(val, other);
            com_m[c + g.nc] = den;
            unsigned long long a = (unsigned long long)(q ? b[2] - b[1] : b[1] - b[0]);
            pairs += a * (a - 1);
        }
    }
    if (force_pairs) block_add_u64(pairs, force_pairs);  // stats only (uniform branch)
}

// stats only: the ordered pairs of the force phase, when k_com (which counts them) did not run
__global__ void k_count_pairs(int ncells, const int *__restrict__ cell_start, unsigned long long *force_pairs) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long pairs = 0;
    if (c < ncells) {
        const unsigned long long a = (unsigned long long)(cell_start[c + 1] - cell_start[c]);
        pairs = a * (a - 1);
    }
    block_add_u64(pairs, force_pairs);
}

// Single strip: the ghost rows are the strip's own last / first row (periodic wrap,
// serial/parsim.cpp:332-333).  Multi-strip runs fill them over NCCL instead.
// It is the first kernel of a single strip's step that every step runs, so it also zeroes the
// step's work-list counters (k_step_reset's job: one launch less per timestep, which is what a
// launch-bound problem such as BASELINE configs[1] is made of).
__global__ void k_ghost_wrap(Geom g, double2 *com_xy, double *com_m, Counters *ctr, int *big_count, int *move_big_count) {
    int gx = blockIdx.x * blockDim.x + threadIdx.x;
    if (gx == 0 && ctr) {
        ctr->n_movers = 0;
        if (big_count) { big_count[0] = 0; big_count[1] = 0; }
        *move_big_count = 0;
    }
    if (gx >= g.nc) return;
    com_xy[gx] = com_xy[g.rows * g.nc + gx];
    com_m[gx] = com_m[g.rows * g.nc + gx];
    com_xy[(g.rows + 1) * g.nc + gx] = com_xy[g.nc + gx];
    com_m[(g.rows + 1) * g.nc + gx] = com_m[g.nc + gx];
}

// ---------------------------------------------------------------------------------------
// Forces + integration, fused (north_star (c)+(d)).
//   own cell : direct pair sum over the cell's particles (serial/parsim.cpp:356-366,:127-148)
//              -- the own cell's centre of mass is never a source (:305)
//   8 others : centre of mass, shifted by +-side across the periodic border (:301-354,:109-125)
//   then applyForce (:150-195): a = f/m, x += v dt + a dt^2/2, v += a dt, fmod wrap, new cell.
// The force on i is G m_i sum_k m_k d_k/|d_k|^3, so a_i = G sum_k m_k d_k/|d_k|^3 needs no
// division by m_i; |d|^-3 comes from one FP64 rsqrt instead of the reference's sqrt + 3
// divisions (ulp-level differences).  Coincident points (d == 0, including
// k == i) contribute nothing, as in :134-137.  The integration itself keeps the reference's
// rounding sequence exactly (no FMA contraction).
// Pre-move positions are read from pos_in by every thread of the cell, so post-move
// positions go to pos_out.
// One source (particle or neighbour centre of mass) acting on a target: a += m_k d / |d|^3.
// |d|^-3 comes from the MUFU.RSQ64H seed y (rel. error < 2^-21) and ONE third-order correction
// of the cube: with e = 1 - d2 y^2,  d2^(-3/2) = y^3 (1 + e (3/2 + 15/8 e)) + O(e^3) -- 13 FP64
// instructions per pair where the reference's sqrt + 3 divisions cost ~60.  kTinyD2 keeps
// coincident points (d == 0, including k == i) finite: they contribute exactly 0 like the
// reference's `if (d == 0) return` (:134-137); it is far below the square of any position ulp.
constexpr double kTinyD2 = 1e-60;

__device__ __forceinline__ void accumulate(double dx, double dy, double mk, double &ax, double &ay) {
    const double d2 = fma(dx, dx, fma(dy, dy, kTinyD2));
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d2));
    const double yy = y * y;
    const double e = fma(-d2, yy, 1.0);
    const double p = fma(e, 1.875, 1.5);
    const double u = (mk * y) * yy;
    const double w = fma(u * e, p, u);
    ax = fma(w, dx, ax);
    ay = fma(w, dy, ay);
}

// x = fmod(x + side, side) (:171-172) without the library loop: fmod is exact, and for
// t = x + side in [side, 4 side) the subtractions below are exact too (Sterbenz), so the
// result is bit-identical.  Anything else (a jump of more than a box length) takes fmod.
__device__ __forceinline__ double wrap_coord(double x, double side) {
    const double t = __dadd_rn(x, side);
    if (t >= side) {
        const double r1 = __dadd_rn(t, -side);
        if (r1 < side) return r1;
        const double r2 = __dadd_rn(r1, -side);  // r1 in [side, 3 side): exact again
        if (r2 < side && t < 4.0 * side) return r2;
        return fmod(t, side);
    }
    return t >= 0.0 ? t : fmod(t, side);
}

// cell_coord without the division: q = v * (1/cellw) differs from the correctly rounded
// v / cellw by < 4 ulp, so trunc() agrees unless q sits within 1e-12 q of an integer; only
// then the true division of the reference decides.
__device__ __forceinline__ int cell_coord_fast(double v, const Geom &g) {
    const double q = v * g.inv_cellw;
    int c = __double2int_rz(q);
    const double fr = q - (double)c;
    const double tol = q * 1e-12;
    if (fr < tol || fr > 1.0 - tol || !(q >= 0.0)) c = __double2int_rz(__ddiv_rn(v, g.cellw));
    return min(max(c, 0), g.nc - 1);
}
__device__ __forceinline__ int global_cell_fast(double2 p, const Geom &g) {
    return cell_coord_fast(p.y, g) * g.nc + cell_coord_fast(p.x, g);
}

// Pair-interaction microbenchmark: accumulate() on register operands only (no memory traffic),
// four independent targets per thread.  Its rate is the ceiling of the force kernel's inner loop
// for THIS instruction mix (13 FP64 + 1 MUFU.RSQ64H per pair), which is lower than the FMA-chain
// peak because the MUFU shares the issue port of the FP64 pipe.
__global__ void k_pair_peak(int iters, double *out) {
    double ax0 = 0, ay0 = 0, ax1 = 0, ay1 = 0, ax2 = 0, ay2 = 0, ax3 = 0, ay3 = 0;
    double qx = 1.0 + threadIdx.x * 1e-3, qy = 2.0 + blockIdx.x * 1e-6;
    const double px0 = 0.1, px1 = 0.2, px2 = 0.3, px3 = 0.4, py = 0.5, m = 1.25;
    for (int i = 0; i < iters; i++) {
        accumulate(qx - px0, qy - py, m, ax0, ay0);
        accumulate(qx - px1, qy - py, m, ax1, ay1);
        accumulate(qx - px2, qy - py, m, ax2, ay2);
        accumulate(qx - px3, qy - py, m, ax3, ay3);
        qx += 1e-3;  // a new source every iteration
        qy -= 1e-3;
    }
    const double r = ax0 + ay0 + ax1 + ay1 + ax2 + ay2 + ax3 + ay3;
    if (r == 123.456) out[0] = r;
}

// Particle::updatePositionAndVelocity (serial/parsim.cpp:54-60, same rounding sequence), the wrap
// (:171-172), the new cell (:174-177) and the mover list: the tail of a particle's force stage.
__device__ __forceinline__ void integrate_store(const Geom &g, int i, int c, double2 p, double2 v, double ax, double ay,
                                                double2 *__restrict__ pos_out, float2 *__restrict__ posf_out,
                                                double2 *__restrict__ vel, int *__restrict__ newcell,
                                                unsigned char *__restrict__ flags, int *__restrict__ mover_list,
                                                int *n_movers) {
    double x = __dadd_rn(p.x, __dadd_rn(__dmul_rn(v.x, kDeltaT),
                                        __dmul_rn(__dmul_rn(__dmul_rn(0.5, ax), kDeltaT), kDeltaT)));
    double y = __dadd_rn(p.y, __dadd_rn(__dmul_rn(v.y, kDeltaT),
                                        __dmul_rn(__dmul_rn(__dmul_rn(0.5, ay), kDeltaT), kDeltaT)));
    v.x = __dadd_rn(v.x, __dmul_rn(ax, kDeltaT));
    v.y = __dadd_rn(v.y, __dmul_rn(ay, kDeltaT));
    x = wrap_coord(x, g.side);  // :171-172
    y = wrap_coord(y, g.side);
    double2 pn = make_double2(x, y);
    pos_out[i] = pn;
    if (posf_out) posf_out[i] = make_float2((float)x, (float)y);  // pre-filter of k_collide_find
    vel[i] = v;
    const int cnew = global_cell_fast(pn, g);  // :174-177
    newcell[i] = cnew;
    // movers (:183-190) are rare: they go on a list for k_classify, everybody else is done here
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

// A block of TB consecutive particles touches the segments of the cells from its first to its
// last particle's cell: one contiguous range of the (cell, id)-ordered arrays.  The range is
// staged in shared memory with coalesced loads (24 B per source), and the pair loop reads it
// from there: no global load -- and no load latency -- inside the FP64-bound loop.  Ranges
// longer than TILE sources (very full cells) are read through L1 instead.
// TILE = staged sources per block, CELLS = cells per block whose 8 neighbour centres of mass are
// staged (0: read through L1 per particle), MINB = resident blocks per SM the register budget is
// set for: full cells want the deep unrolled pair loop (many registers), sparse cells want
// occupancy (the kernel is then a chain of dependent loads per particle).
// 16- / 8-byte asynchronous copies global -> shared (LDGSTS): the data never passes through
// registers, so the issuing thread does not wait for it until cp_async_wait_all().
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// EARLY: the sources are staged with asynchronous copies and the eight neighbour centres of mass
// are applied (read through L1) BEFORE the block waits for the staged range -- the barrier and the
// load latency in front of the pair loop (15 % + 13 % of the warp stall samples of the plain kernel,
// profiles/ncu_force_r02_v2_stalls.txt) overlap with 104 FP64 instructions per thread.  Same
// operations in the same order per target: bit-identical results.
// ACC_ONLY: the kernel stops after the force sum and stores the acceleration (dbg_acc is the output);
// k_integrate_acc finishes the step once the velocities are there (uploads with late velocities).
template <int TB, int UNROLL, int TILE, int CELLS, int MINB, bool EARLY = false, bool ACC_ONLY = false>
__global__ void __launch_bounds__(TB, MINB)
k_force_integrate(Geom g, int n_val, const int *__restrict__ n_ptr, const double2 *__restrict__ pos_in,
                  const double *__restrict__ m, const int *__restrict__ cell,
                  const int *__restrict__ cell_start, const double2 *__restrict__ com_xy,
                  const double *__restrict__ com_m, double2 *__restrict__ pos_out,
                  float2 *__restrict__ posf_out, double2 *__restrict__ vel, int *__restrict__ newcell,
                  unsigned char *__restrict__ flags, int *__restrict__ mover_list, int *n_movers,
                  double2 *__restrict__ dbg_acc) {
    __shared__ double2 s_pos[TILE];
    __shared__ double s_m[TILE];
    __shared__ double2 s_nxy[CELLS > 0 ? CELLS * 8 : 1];
    __shared__ double s_nm[CELLS > 0 ? CELLS * 8 : 1];
    // deferred steps (small problems): the count comes from device memory and the grid is sized for an
    // upper bound; otherwise the host knows it (the extra dependent load costs latency-bound kernels
    // up to a third of their time at 1e8 particles, profiles/bench_r02_v5_config4w.json)
    const int n = n_ptr ? *n_ptr : n_val;
    const int b0 = blockIdx.x * TB;
    if (b0 >= n) return;
    const int i = b0 + threadIdx.x;
    const int last = min(b0 + TB, n) - 1;
    const int r0 = cell_start[cell[b0]], r1 = cell_start[cell[last] + 1];
    const bool staged = r1 - r0 <= TILE;  // block-uniform
    int c = 0, s = 0, e = 0;
    double2 p = make_double2(0.0, 0.0);
    if (i < n) {
        c = cell[i];
        p = pos_in[i];
        s = cell_start[c];
        e = cell_start[c + 1];
    }
    if (staged) {
        for (int k = r0 + threadIdx.x; k < r1; k += TB) {
            if (EARLY) {
                cp_async16(&s_pos[k - r0], &pos_in[k]);
                cp_async8(&s_m[k - r0], &m[k]);
            } else {
                s_pos[k - r0] = pos_in[k];
                s_m[k - r0] = m[k];
            }
        }
    }
    // The 3 x 3 neighbourhood (north_star (c)): the 8 neighbour centres of mass of every cell this
    // block touches are staged in shared memory ONCE per cell, already shifted by +-side across
    // the periodic border (:314-329) -- the wrap arithmetic is per (cell, neighbour), not per
    // (particle, neighbour), and the particles of a cell read the same 8 entries (broadcast).
    const int cf = cell[b0], ncb = cell[last] - cf + 1;
    const bool nb_staged = !EARLY && CELLS > 0 && ncb <= CELLS;  // block-uniform
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
    double ax = 0.0, ay = 0.0;
    double2 v = make_double2(0.0, 0.0);
    if (!ACC_ONLY && i < n) v = vel[i];
    if (!EARLY) {
        if (staged || nb_staged) __syncthreads();
        if (i >= n) return;
    }
    if (EARLY && i >= n) {
        // nothing to compute, but the block still meets at the barrier below
    } else if (nb_staged) {
        const int base = (c - cf) * 8;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const double2 cm = s_nxy[base + j];
            accumulate(cm.x - p.x, cm.y - p.y, s_nm[base + j], ax, ay);
        }
    } else {
        const int lrow = c / g.nc;
        const int gx = c - lrow * g.nc;
        const int grow = g.row0 + lrow;
        if ((unsigned)(gx - 1) < (unsigned)(g.nc - 2) && (unsigned)(grow - 1) < (unsigned)(g.nc - 2)) {
            // interior cell: no periodic shift, the eight neighbours sit at fixed offsets (same
            // order as the general loop below: dx outer, dy inner)
            const int base = c + g.nc;  // + the ghost row
            const int off[8] = {-g.nc - 1, -1, g.nc - 1, -g.nc, g.nc, -g.nc + 1, 1, g.nc + 1};
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const double2 cm = __ldg(&com_xy[base + off[j]]);
                const double M = __ldg(&com_m[base + off[j]]);
                accumulate(cm.x - p.x, cm.y - p.y, M, ax, ay);
            }
        } else
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
    if (EARLY) {
        if (staged) {
            cp_async_wait_all();
            __syncthreads();
        }
        if (i >= n) return;
    }
    if (staged) {
#pragma unroll UNROLL
        for (int k = s - r0; k < e - r0; k++) {
            const double2 q = s_pos[k];
            accumulate(q.x - p.x, q.y - p.y, s_m[k], ax, ay);
        }
    } else {
#pragma unroll 4
        for (int k = s; k < e; k++) {
            const double2 q = __ldg(&pos_in[k]);
            accumulate(q.x - p.x, q.y - p.y, __ldg(&m[k]), ax, ay);
        }
    }
    ax *= kG;
    ay *= kG;
    if (ACC_ONLY) {
        dbg_acc[i] = make_double2(ax, ay);
        return;
    }
    if (dbg_acc) dbg_acc[i] = make_double2(ax, ay);
    integrate_store(g, i, c, p, v, ax, ay, pos_out, posf_out, vel, newcell, flags, mover_list, n_movers);
}

// The second half of a step whose force sums were taken early (k_force_integrate<.., ACC_ONLY>).
__global__ void __launch_bounds__(kThreads)
k_integrate_acc(Geom g, int n, const double2 *__restrict__ pos_in, const int *__restrict__ cell,
                const double2 *__restrict__ acc, double2 *__restrict__ pos_out, float2 *__restrict__ posf_out,
                double2 *__restrict__ vel, int *__restrict__ newcell, unsigned char *__restrict__ flags,
                int *__restrict__ mover_list, int *n_movers) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double2 a = acc[i];
    integrate_store(g, i, cell[i], pos_in[i], vel[i], a.x, a.y, pos_out, posf_out, vel, newcell, flags, mover_list, n_movers);
}

// ---------------------------------------------------------------------------------------
// List update (updateCellParticles, serial/parsim.cpp:234-259).  The reference erases at the
// loop position and still advances it (:251-253): inside a run of consecutive leavers of a
// cell's bin-time list, run positions 0,2,4,... are moved to their new cell's list and
// 1,3,5,... stay listed in the old cell for this step's collision phase ("strays").
// The bin-time list is id-ordered and contains dead placeholders, which break runs; the
// alive part of it is exactly this cell's segment, the dead part is the tombstone CSR.
__device__ __forceinline__ bool tombstone_between(const int *__restrict__ dead_start,
                                                  const unsigned *__restrict__ dead_id, int c,
                                                  unsigned lo, unsigned hi) {
    if (!dead_start) return false;
    for (int e = dead_start[c], end = dead_start[c + 1]; e < end; e++) {
        unsigned d = dead_id[e];
        if (d > lo && d < hi) return true;
    }
    return false;
}

__global__ void __launch_bounds__(kThreads)
k_classify(Geom g, const int *__restrict__ mover_list, const int *__restrict__ n_movers,
           const int *__restrict__ cell, const int *__restrict__ cell_start,
           const int *__restrict__ newcell, const unsigned *__restrict__ id,
           const int *__restrict__ dead_start, const unsigned *__restrict__ dead_id,
           unsigned char *__restrict__ flags, int *arr_head, int *__restrict__ arr_next,
           const double2 *__restrict__ pos_new, int *arr_cnt, ArrRec *__restrict__ arr_rec,
           unsigned long long *stat_strays, int *__restrict__ emig_list, int *emig_count, int emig_cap,
           Counters *ctr) {
    const int nm = *n_movers;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < nm; t += gridDim.x * blockDim.x) {
        const int i = mover_list[t];
        const int c = cell[i];
        const int cg = c + g.cell_base;
        const int dstg = newcell[i];
        // length of the run of consecutive leavers that ends just before i
        int run = 0;
        const int s = g.no_strays ? i : cell_start[c];
        for (int u = i - 1; u >= s; u--) {
            if (newcell[u] == cg) break;
            if (tombstone_between(dead_start, dead_id, c, id[u], id[u + 1])) break;
            run++;
        }
        unsigned char f = F_MOVER;
        if ((run & 1) == 0) {
            f |= F_REMOVED;
            int dst = dstg - g.cell_base;
            if (dst >= 0 && dst < g.ncells) {
                arr_next[i] = atomicExch(&arr_head[dst], i);  // unordered bag; order comes from keys
                if (arr_cnt) {
                    const int sl = atomicAdd(&arr_cnt[dst], 1);
                    if (sl < kArrSlots) {
                        ArrRec r;
                        r.pos = pos_new[i];
                        r.key = (1ull << 63) | ((unsigned long long)(unsigned)cg << 32) | id[i];
                        r.idx = i;
                        r.pad = 0;
                        arr_rec[(size_t)dst * kArrSlots + sl] = r;
                    }
                }
            } else {
                f |= F_EMIGRANT;  // multi-strip only: packed for the exchange from this short list
                const int sl = atomicAdd(emig_count, 1);
                if (sl < emig_cap) emig_list[sl] = i;
                else ctr->error = DE_MIGRANT_OVERFLOW;
            }
        } else {
            atomicAdd(stat_strays, 1ull);  // rare
        }
        flags[i] = f;
    }
}

// ---------------------------------------------------------------------------------------
// Collisions (checkCollisions, serial/parsim.cpp:388-420).  The reference sweeps each cell's
// post-move list greedily (j ascending, k > j ascending) and counts a close pair iff neither
// member is in the cell's set yet (:404-409); every member of any close pair dies (:414-418).
// Equivalent order-free rule: with first(j) = the earliest list entry that is close to j,
// a collision is counted exactly for the pairs with first(j) = k and first(k) = j.
// "Earliest" is decided by a 64-bit list key instead of a memory position:
//   stayers and strays keep their bin-time place  -> key = id
//   arrivals are appended by (source cell, id)    -> key = 1<<63 | source cell << 32 | id
// close(j,k) <=> sqrt(dx*dx + dy*dy) < EPSILON (:76-79,:400) <=> dx*dx + dy*dy < thr2 where
// thr2 is the smallest double whose correctly rounded sqrt is >= EPSILON (host-computed), and
// the sum is rounded exactly like the reference's (no FMA).
__device__ __forceinline__ unsigned long long list_key(unsigned char f, int src_cell_global,
                                                       unsigned id) {
    if (f & F_REMOVED)
        return (1ull << 63) | ((unsigned long long)(unsigned)src_cell_global << 32) | id;
    return id;
}

__device__ __forceinline__ bool is_close(double2 a, double2 b, double thr2) {
    double dx = a.x - b.x, dy = a.y - b.y;
    return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)) < thr2;
}

// The first list entry (by list key) that is close to entry i, or -1: first(i) of the mutual-first
// rule.  Entry i's list L is the cell it was binned in, or -- for a removed mover -- the cell it
// moved to.  `tested` counts the candidates looked at (stats).
__device__ __forceinline__ int find_first_partner(const Geom &g, int i, int n_res, double thr2, float thrf,
                                                  const double2 *__restrict__ pos, const float2 *__restrict__ posf,
                                                  const int *__restrict__ cell, const int *__restrict__ cell_start,
                                                  const int *__restrict__ newcell, const unsigned *__restrict__ id,
                                                  const unsigned char *__restrict__ flags,
                                                  const int *__restrict__ arr_head, const int *__restrict__ arr_next,
                                                  const int *__restrict__ guest_src, unsigned long long &tested) {
    const unsigned char f = flags[i];
    int L = cell[i];  // issued together with flags[i]; only the rare removed mover pays a second, dependent load
    int best_idx = -1;
    if (f & F_EMIGRANT) return -1;  // emigrants are checked on the strip they moved to
    if (f & F_REMOVED) L = newcell[i] - g.cell_base;
    const double2 p = pos[i];
    const float2 pf = posf[i];
    unsigned long long best = ~0ull;
    const int s = cell_start[L], e = cell_start[L + 1];
    int k_arr = arr_head[L];  // loaded with the bounds, not after the loop: one dependent round trip less
    // residents of L that are still listed there (stayers and strays), key = id, id-sorted.
    // FP32 pre-filter: thrf over-covers the rounding of the float copies, so every pair
    // that is close in FP64 passes it; the decision itself is the exact FP64 test.
#pragma unroll 4
    for (int k = s; k < e; k++) {
        float2 qf = __ldg(&posf[k]);
        float fx = qf.x - pf.x, fy = qf.y - pf.y;
        if (fmaf(fx, fx, fy * fy) < thrf) {
            if (k != i && !(flags[k] & F_REMOVED) && is_close(p, pos[k], thr2)) {
                unsigned long long key = id[k];
                if (key < best) { best = key; best_idx = k; }
            }
        }
    }
    // arrivals appended to L's list
    int walked = 0;
    for (int k = k_arr; k >= 0; k = arr_next[k]) {
        walked++;
        if (k == i) continue;
        if (is_close(p, pos[k], thr2)) {
            int src = k < n_res ? cell[k] + g.cell_base : guest_src[k - n_res];
            unsigned long long key = list_key(F_REMOVED, src, id[k]);
            if (key < best) { best = key; best_idx = k; }
        }
    }
    tested += (unsigned long long)(e - s + walked);
    return best_idx;
}

// entries = *n_res_ptr residents + n_guest guests that arrived from other strips (those have
// cell[i] = -1 and carry their source cell in guest_src[i - n_res]).
#ifndef PS_FIND_MINB
#define PS_FIND_MINB 8  // 32 registers: the search is a chain of dependent loads, occupancy decides (6 blocks: +35 %, profiles/gpurun_r02_v7.log)
#endif
__global__ void __launch_bounds__(kThreads, PS_FIND_MINB)
k_collide_find(Geom g, int n_res_val, const int *__restrict__ n_res_ptr, int n_guest, double thr2, float thrf,
               const double2 *__restrict__ pos,
               const float2 *__restrict__ posf,
               const int *__restrict__ cell, const int *__restrict__ cell_start,
               const int *__restrict__ newcell, const unsigned *__restrict__ id,
               const unsigned char *__restrict__ flags, const int *__restrict__ arr_head,
               const int *__restrict__ arr_next, const int *__restrict__ guest_src,
               int *__restrict__ partner, unsigned long long *collision_pairs) {
    const int n_res = n_res_ptr ? *n_res_ptr : n_res_val, n = n_res + n_guest;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long tested = 0;
    if (i < n)
        partner[i] = find_first_partner(g, i, n_res, thr2, thrf, pos, posf, cell, cell_start, newcell, id, flags,
                                        arr_head, arr_next, guest_src, tested);
    if (collision_pairs) block_add_u64(tested, collision_pairs);  // stats only (uniform branch)
}

// ---------------------------------------------------------------------------------------
__global__ void k_set_int(int *dst, int v) { *dst = v; }

// Tombstone counts: total per cell += this step's deaths (and the per-step array is cleared).
__global__ void k_add_counts(int n, int *__restrict__ acc, int *__restrict__ add) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) { acc[c] += add[c]; add[c] = 0; }
}

// start of a step: the three work-list counters of the step in one launch (instead of three memsets)
__global__ void k_step_reset(Counters *ctr, int *big_count, int *move_big_count) {
    ctr->n_movers = 0;
    if (big_count) { big_count[0] = 0; big_count[1] = 0; }
    *move_big_count = 0;
}

// ---------------------------------------------------------------------------------------
// Two scans in one pass (the end of a step needs the new cell_start AND the new tombstone_start):
// A is scanned as it is; B is (acc + add), and acc += add, add = 0 is done on the way (the running
// tombstone count per cell takes up this step's deaths).  tile_sums holds 2 * ntiles entries.
__global__ void k_scan2_tile_sums(const int *__restrict__ a, const int *__restrict__ acc, const int *__restrict__ add,
                                  int N, int ntiles, int *tile_sums) {
    __shared__ int ws[33];
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int sa = 0, sb = 0;
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        if (k < N) { sa += a[k]; sb += acc[k] + add[k]; }
    }
    int tot;
    block_exclusive_scan(sa, ws, tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
    block_exclusive_scan(sb, ws, tot);
    if (threadIdx.x == 0) tile_sums[ntiles + blockIdx.x] = tot;
}

__global__ void k_scan2_top(int *tile_sums, int ntiles, int *total_a, int *total_a2, int *total_b) {
    __shared__ int ws[33];
    for (int half = 0; half < 2; half++) {
        int *ts = tile_sums + half * ntiles;
        int carry = 0;
        for (int base = 0; base < ntiles; base += kThreads) {
            int k = base + threadIdx.x;
            int v = k < ntiles ? ts[k] : 0;
            int tot;
            int ex = block_exclusive_scan(v, ws, tot);
            if (k < ntiles) ts[k] = carry + ex;
            carry += tot;
        }
        if (threadIdx.x == 0) {
            if (half == 0) { *total_a = carry; if (total_a2) *total_a2 = carry; }
            else *total_b = carry;
        }
        __syncthreads();
    }
}

// Grids of up to kScanFewTiles tiles: every block of the apply kernel adds up the tile sums before
// its own by itself, so the one-block k_scan2_top launch in the middle disappears (two launches
// instead of three: launch-bound problems such as BASELINE configs[1] are made of launches).  A
// one-block scan of the whole grid was measured SLOWER than the three launches (30-36 us against
// 20 us at 1e4 cells, with 256 and with 1024 threads: one SM's load/store path is the limit).
constexpr int kScanFewTiles = kThreads;
__global__ void k_scan2_apply_few(const int *__restrict__ a, int *__restrict__ acc, int *__restrict__ add, int N, int ntiles,
                                  const int *__restrict__ tile_sums, int *__restrict__ out_a, int *total_a2,
                                  int *__restrict__ out_b) {
    __shared__ int ws[33];
    // tile sums before mine (and, for the last block, the totals)
    const int ta = (int)threadIdx.x < ntiles ? tile_sums[threadIdx.x] : 0;
    const int tb = (int)threadIdx.x < ntiles ? tile_sums[ntiles + threadIdx.x] : 0;
    int tot_a, tot_b, tot;
    block_exclusive_scan((int)threadIdx.x < (int)blockIdx.x ? ta : 0, ws, tot_a);
    block_exclusive_scan((int)threadIdx.x < (int)blockIdx.x ? tb : 0, ws, tot_b);
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int va[kScanItems], vb[kScanItems];
    int sa = 0, sb = 0;
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        va[j] = k < N ? a[k] : 0;
        vb[j] = k < N ? acc[k] + add[k] : 0;
        sa += va[j];
        sb += vb[j];
    }
    int ea = block_exclusive_scan(sa, ws, tot) + tot_a;
    const int mine_a = tot;
    int eb = block_exclusive_scan(sb, ws, tot) + tot_b;
    const int mine_b = tot;
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        if (k < N) {
            out_a[k] = ea;
            out_b[k] = eb;
            acc[k] = vb[j];
            add[k] = 0;
        }
        ea += va[j];
        eb += vb[j];
    }
    if ((int)blockIdx.x == ntiles - 1 && threadIdx.x == 0) {
        out_a[N] = tot_a + mine_a;
        if (total_a2) *total_a2 = tot_a + mine_a;
        out_b[N] = tot_b + mine_b;
    }
}

__global__ void k_scan2_apply(const int *__restrict__ a, int *__restrict__ acc, int *__restrict__ add, int N, int ntiles,
                              const int *__restrict__ tile_sums, int *__restrict__ out_a, int *__restrict__ out_b) {
    __shared__ int ws[33];
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int va[kScanItems], vb[kScanItems];
    int sa = 0, sb = 0;
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        va[j] = k < N ? a[k] : 0;
        vb[j] = k < N ? acc[k] + add[k] : 0;
        sa += va[j];
        sb += vb[j];
    }
    int tot;
    int ea = block_exclusive_scan(sa, ws, tot) + tile_sums[blockIdx.x];
    int eb = block_exclusive_scan(sb, ws, tot) + tile_sums[ntiles + blockIdx.x];
#pragma unroll
    for (int j = 0; j < kScanItems; j++) {
        int k = base + j;
        if (k < N) {
            out_a[k] = ea;
            out_b[k] = eb;
            acc[k] = vb[j];
            add[k] = 0;
        }
        ea += va[j];
        eb += vb[j];
    }
}

// ---------------------------------------------------------------------------------------
// Dumps (tests): scatter the device state into id-ordered arrays.
__global__ void k_dump_alive(Geom g, int n, const double2 *__restrict__ pos,
                             const double2 *__restrict__ vel, const double *__restrict__ m,
                             const unsigned *__restrict__ id, const int *__restrict__ cell,
                             const int *__restrict__ newcell, const unsigned char *__restrict__ flags,
                             int have_new, int have_flags, double *x, double *y, double *vx, double *vy,
                             double *mass, int *alive, int *cell_out, int *list_cell,
                             unsigned long long *list_rank, unsigned char *owned) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned p = id[i];
    double2 a = pos[i], v = vel[i];
    x[p] = a.x; y[p] = a.y; vx[p] = v.x; vy[p] = v.y; mass[p] = m[i];
    alive[p] = 1;
    owned[p] = 1;
    int cg = cell[i] + g.cell_base;
    cell_out[p] = have_new ? newcell[i] : cg;
    unsigned char f = have_flags ? flags[i] : 0;
    list_cell[p] = (f & F_REMOVED) ? newcell[i] : cg;
    list_rank[p] = list_key(f, cg, p);
}

__global__ void k_dump_dead(int n, const double2 *__restrict__ pos, const double2 *__restrict__ vel,
                            const unsigned *__restrict__ id, const int *__restrict__ cellg, double *x,
                            double *y, double *vx, double *vy, double *mass, int *alive, int *cell_out,
                            int *list_cell, unsigned long long *list_rank, unsigned char *owned) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned p = id[i];
    double2 a = pos[i], v = vel[i];
    x[p] = a.x; y[p] = a.y; vx[p] = v.x; vy[p] = v.y; mass[p] = 0.0;
    alive[p] = 0;
    owned[p] = 1;
    cell_out[p] = cellg[i];
    list_cell[p] = -1;
    list_rank[p] = 0;
}

// particle 0 for parsim_result: alive arrays first, then the graveyard
__global__ void k_find_id(int n, const unsigned *__restrict__ id, const double2 *__restrict__ pos,
                          unsigned want, double2 *out, int *found) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (id[i] == want) { *out = pos[i]; *found = 1; }
}

}  // namespace parsim
