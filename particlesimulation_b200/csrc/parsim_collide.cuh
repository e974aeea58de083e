// parsim_collide.cuh -- cell-centric collision search (checkCollisions, serial/parsim.cpp:388-420).
//
// k_collide_find (parsim_kernels.cuh) tests every list entry against every other entry of its
// cell: O(n_c^2) distance tests.  EPSILON is usually far smaller than a cell, so here each cell's
// post-move list is first binned into an nb x nb sub-grid whose pitch h >= EPSILON (strays that lie
// outside the cell are clamped to the border bins: clamping never increases index distances), and
// an entry is tested only against the 3 x 3 bins around its own -- close(p, q) implies adjacent
// bins.  The decision itself is the same exact FP64 predicate, and "first close partner" is the
// minimum list key over the candidates, so the result is identical to the all-pairs sweep.
//
//   k_collide_cells : one WARP per cell, the cell's list staged in shared memory in bin order;
//                     cells whose list exceeds the staging capacity go to a worklist
//   k_collide_big   : one BLOCK per worklist cell (large staging area, 32 x 32 bins at most);
//                     lists that do not fit even there are swept all-pairs
#pragma once
#include "parsim_kernels.cuh"

namespace parsim {

struct CollideArgs {
    Geom g;
    int n_res_val;        // residents [0, n_res), guests from other strips behind them;
    const int *n_res_ptr; // n_res = *n_res_ptr when non-null (deferred steps), else n_res_val
    double thr2;          // squared-distance threshold (see k_collide_find)
    const double2 *pos;   // post-move positions
    const int *cell, *cell_start, *newcell;
    const unsigned *id;
    const unsigned char *flags;
    const int *arr_head, *arr_next, *guest_src;
    int *partner;
    unsigned long long *collision_pairs;  // stats (nullptr: off)
    int nb;               // sub-grid bins per side (>= 1)
    double inv_h;         // bins per unit length
    int cap;              // staging capacity (list entries) of one group
    int *worklist, *work_count;
    const int *arr_cnt;   // arrival records (k_classify): count per cell, kArrSlots records per cell
    const ArrRec *arr_rec;
};

__device__ __forceinline__ unsigned long long entry_key(const CollideArgs &a, int k) {
    if (a.flags[k] & F_REMOVED) {  // an arrival: appended after the bin-time entries, by (source cell, id)
        const int n_res = a.n_res_ptr ? *a.n_res_ptr : a.n_res_val;
        const int src = k < n_res ? a.cell[k] + a.g.cell_base : a.guest_src[k - n_res];
        return list_key(F_REMOVED, src, a.id[k]);
    }
    return a.id[k];
}

__device__ __forceinline__ int sub_bin(const CollideArgs &a, double2 p, double x0, double y0) {
    int bx = __double2int_rd((p.x - x0) * a.inv_h);
    int by = __double2int_rd((p.y - y0) * a.inv_h);
    bx = min(max(bx, 0), a.nb - 1);
    by = min(max(by, 0), a.nb - 1);
    return by * a.nb + bx;
}

// warp kernel: rows of the sub-grid are always 16 bins apart (a.nb <= 16; the unused bins stay
// empty), so row and column of a bin are a shift and a mask instead of a division
__device__ __forceinline__ int sub_bin16(const CollideArgs &a, double2 p, double x0, double y0) {
    int bx = __double2int_rd((p.x - x0) * a.inv_h);
    int by = __double2int_rd((p.y - y0) * a.inv_h);
    bx = min(max(bx, 0), a.nb - 1);
    by = min(max(by, 0), a.nb - 1);
    return (by << 4) | bx;
}

template <int G>
__device__ __forceinline__ void group_sync() {
    if (G == 32) __syncwarp(); else __syncthreads();
}

// exclusive scan of v[0..n) in place by a group of G threads; v[n] = total.  n <= 1024.
template <int G>
__device__ __forceinline__ void group_scan(int *v, int n, int t, int *scratch) {
    const int per = (n + G - 1) / G;
    const int lo = min(t * per, n), hi = min(lo + per, n);
    int s = 0;
    for (int k = lo; k < hi; k++) s += v[k];
    int inc = s;
    const int lane = t & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    int base = inc - s;
    if (G > 32) {
        if (lane == 31) scratch[t >> 5] = inc;
        __syncthreads();
        int carry = 0;
        for (int w = 0; w < (t >> 5); w++) carry += scratch[w];
        base += carry;
        __syncthreads();
    }
    for (int k = lo; k < hi; k++) { int c = v[k]; v[k] = base; base += c; }
    if (t == G - 1) v[n] = base;
    group_sync<G>();
}

// Shared-memory staging area of one group, carved from `raw`:
//   start[nbins + 1], fill[nbins], spos[cap], skey[cap], sidx[cap], sbin[cap]
struct Staging {
    int *start, *fill, *sidx;
    double2 *spos;
    unsigned long long *skey;
    unsigned short *sbin;
};
__host__ __device__ inline size_t staging_bytes(int nbins, int cap) {
    size_t b = sizeof(int) * (size_t)(2 * nbins + 2);
    b = (b + 15) & ~(size_t)15;
    return b + (size_t)cap * (16 + 8 + 4 + 2) + 16;
}
__device__ __forceinline__ Staging carve(unsigned char *raw, int nbins, int cap) {
    Staging s;
    s.start = (int *)raw;
    s.fill = s.start + nbins + 1;
    size_t off = (sizeof(int) * (size_t)(2 * nbins + 2) + 15) & ~(size_t)15;
    s.spos = (double2 *)(raw + off);
    s.skey = (unsigned long long *)(s.spos + cap);
    s.sidx = (int *)(s.skey + cap);
    s.sbin = (unsigned short *)(s.sidx + cap);
    return s;
}

// The list of cell L: bin-time entries that were not moved away (stayers and strays), then the
// arrivals bag.  Returns: >= 0 number of candidate tests done; -1 = list larger than a.cap.
template <int G>
__device__ long long collide_cell(const CollideArgs &a, int L, int t, Staging sm, int *scratch) {
    const int nbins = a.nb * a.nb;
    const int s = a.cell_start[L], e = a.cell_start[L + 1];
    const int head = a.arr_head[L];
    const int lrow = L / a.g.nc, gx = L - lrow * a.g.nc;
    const double x0 = gx * a.g.cellw, y0 = (a.g.row0 + lrow) * a.g.cellw;

    for (int b = t; b < nbins; b += G) { sm.start[b] = 0; sm.fill[b] = 0; }
    group_sync<G>();
    // pass 1: count per bin
    for (int k = s + t; k < e; k += G) {
        const unsigned char f = a.flags[k];
        if (f & F_REMOVED) {
            if (f & F_EMIGRANT) a.partner[k] = -1;  // checked on the strip it moved to
            continue;
        }
        atomicAdd(&sm.start[sub_bin(a, a.pos[k], x0, y0)], 1);
    }
    {
        int j = 0;
        for (int x = head; x >= 0; x = a.arr_next[x], j++)
            if ((j % G) == t) atomicAdd(&sm.start[sub_bin(a, a.pos[x], x0, y0)], 1);
    }
    group_sync<G>();
    group_scan<G>(sm.start, nbins, t, scratch);
    const int total = sm.start[nbins];
    if (total > a.cap) return -1;
    // pass 2: stage in bin order
    for (int k = s + t; k < e; k += G) {
        if (a.flags[k] & F_REMOVED) continue;
        const double2 p = a.pos[k];
        const int b = sub_bin(a, p, x0, y0);
        const int o = sm.start[b] + atomicAdd(&sm.fill[b], 1);
        sm.spos[o] = p;
        sm.skey[o] = a.id[k];
        sm.sidx[o] = k;
        sm.sbin[o] = (unsigned short)b;
    }
    {
        int j = 0;
        for (int x = head; x >= 0; x = a.arr_next[x], j++) {
            if ((j % G) != t) continue;
            const double2 p = a.pos[x];
            const int b = sub_bin(a, p, x0, y0);
            const int o = sm.start[b] + atomicAdd(&sm.fill[b], 1);
            sm.spos[o] = p;
            sm.skey[o] = entry_key(a, x);
            sm.sidx[o] = x;
            sm.sbin[o] = (unsigned short)b;
        }
    }
    group_sync<G>();
    // query: 3 x 3 bins around the entry's own; each bin row is one contiguous range
    long long tested = 0;
    for (int o = t; o < total; o += G) {
        const double2 p = sm.spos[o];
        const int b = sm.sbin[o];
        const int by = b / a.nb, bx = b - by * a.nb;
        const int c0 = max(bx - 1, 0), c1 = min(bx + 1, a.nb - 1);
        unsigned long long best = ~0ull;
        int best_idx = -1;
        for (int r = max(by - 1, 0); r <= min(by + 1, a.nb - 1); r++) {
            const int q0 = sm.start[r * a.nb + c0], q1 = sm.start[r * a.nb + c1 + 1];
            tested += q1 - q0;
            for (int q = q0; q < q1; q++) {
                if (q != o && is_close(p, sm.spos[q], a.thr2)) {
                    const unsigned long long key = sm.skey[q];
                    if (key < best) { best = key; best_idx = sm.sidx[q]; }
                }
            }
        }
        a.partner[sm.sidx[o]] = best_idx;
    }
    group_sync<G>();  // the staging area is reused by the group's next cell
    return tested;
}

// All-pairs sweep for lists that fit no staging area (tens of thousands of entries per cell).
// Such a cell is cut into slices of kThreads bin-time entries (+ one slice for the arrivals bag),
// and k_collide_huge spreads the slices of all such cells over the whole GPU.
// The bin-time entries are id-sorted and every arrival's key is above every id, so the FIRST
// close bin-time entry met in ascending order is the partner (first(i) of the mutual-first rule):
// the sweep stops there.  At the densities that produce such cells a particle has hundreds of
// close neighbours, so a thread looks at a few dozen entries, not at the whole cell.
__device__ __forceinline__ long long sweep_allpairs(const CollideArgs &a, int i, int s, int e, int head) {
    const double2 p = a.pos[i];
    int best_idx = -1;
    int k = s;
    for (; k < e; k++) {
        if (is_close(p, __ldg(&a.pos[k]), a.thr2) && k != i && !(a.flags[k] & F_REMOVED)) {
            best_idx = k;
            break;
        }
    }
    long long walked = k - s;
    if (best_idx < 0) {
        unsigned long long best = ~0ull;
        for (int x = head; x >= 0; x = a.arr_next[x], walked++) {
            if (x != i && is_close(p, a.pos[x], a.thr2)) {
                const unsigned long long key = entry_key(a, x);
                if (key < best) { best = key; best_idx = x; }
            }
        }
    }
    a.partner[i] = best_idx;
    return walked;
}

__global__ void __launch_bounds__(kThreads)
k_collide_huge(CollideArgs a, const int2 *__restrict__ items, const int *__restrict__ n_items) {
    long long tested = 0;
    const int n = *n_items;
    for (int w = blockIdx.x; w < n; w += gridDim.x) {
        const int L = items[w].x, slice = items[w].y;
        const int s = a.cell_start[L], e = a.cell_start[L + 1];
        const int head = a.arr_head[L];
        if (slice >= 0) {  // kThreads bin-time entries
            const int k = s + slice * kThreads + threadIdx.x;
            if (k < e) {
                const unsigned char f = a.flags[k];
                if (f & F_REMOVED) {
                    if (f & F_EMIGRANT) a.partner[k] = -1;
                } else {
                    tested += sweep_allpairs(a, k, s, e, head);
                }
            }
        } else {  // the arrivals bag
            int j = 0;
            for (int x = head; x >= 0; x = a.arr_next[x], j++)
                if ((j % kThreads) == threadIdx.x) tested += sweep_allpairs(a, x, s, e, head);
        }
    }
    if (a.collision_pairs) block_add_u64((unsigned long long)tested, a.collision_pairs);
}

constexpr int kCollideWarps = 8;  // cells per block in k_collide_cells

// Warp-per-cell staging area: start[256 + 8] (always 16 x 16 bins worth), arr_idx[32], spos[cap],
// sord[cap], best[cap]
constexpr int kWarpBins = 256;
__host__ __device__ inline size_t warp_staging_bytes(int cap) {
    return sizeof(int) * (size_t)(kWarpBins + 8 + 32) + (size_t)cap * (16 + 4 + 4);
}

// exclusive scan of cnt[0..256) in place by one warp; cnt[256] = total.  Lane l owns the eight
// consecutive bins 8l..8l+7 (two 16-byte accesses each way).
__device__ __forceinline__ void warp_scan_bins(int *cnt, int lane) {
    int4 *v = reinterpret_cast<int4 *>(cnt) + 2 * lane;
    int4 a = v[0], b = v[1];
    const int s0 = a.x, s1 = s0 + a.y, s2 = s1 + a.z, s3 = s2 + a.w;
    const int s4 = s3 + b.x, s5 = s4 + b.y, s6 = s5 + b.z, s7 = s6 + b.w;
    int inc = s7;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    const int base = inc - s7;
    v[0] = make_int4(base, base + s0, base + s1, base + s2);
    v[1] = make_int4(base + s3, base + s4, base + s5, base + s6);
    if (lane == 31) cnt[kWarpBins] = inc;
    __syncwarp();
}

// One warp per cell, IT bin-time entries per lane held in registers (+ one arrival per lane):
// every global load of the cell is issued before the first one is consumed, so a cell costs
// about one memory latency instead of one per loop iteration.  The smem atomic that counts a
// bin also hands out the entry's rank inside it, so the list is walked once.
// List order is a 32-bit value here: bin-time entries are already in key order (the segment is
// id-sorted), so ord = position in the segment; arrivals follow, ordered by (source cell, id).
// Each unordered candidate pair is tested ONCE (own bin behind the entry, the next bin of the
// row, three bins of the next row) and a close pair lowers `best` of both members (smem min).
#ifndef PS_COLLIDE_MINB
#define PS_COLLIDE_MINB 5  // 48 registers; 1 (more registers) and 7-8 (spills) are both 15 % slower
#endif
template <int IT>
__global__ void __launch_bounds__(kCollideWarps * 32, PS_COLLIDE_MINB)
k_collide_cells(CollideArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int L = blockIdx.x * kCollideWarps + warp;
    constexpr int cap = 32 * IT + 32;
    unsigned char *raw = smem_raw + (size_t)warp * warp_staging_bytes(cap);
    int *start = (int *)raw;
    int *arr_idx = start + kWarpBins + 8;
    double2 *spos = (double2 *)(raw + sizeof(int) * (size_t)(kWarpBins + 8 + 32));
    unsigned *sord = (unsigned *)(spos + cap);
    unsigned *best = sord + cap;
    long long tested = 0;
    if (L < a.g.ncells) {  // warp-uniform
        const int s = a.cell_start[L], e = a.cell_start[L + 1];
        const int n_rec = a.arr_cnt[L];
        if (e - s > 32 * IT) {  // warp-uniform
            if (lane == 0) a.worklist[atomicAdd(a.work_count, 1)] = L;
        } else {
            // every load of the bin-time entries is in flight before the arrivals are looked at
            double2 p[IT + 1];
            unsigned char f[IT];
#pragma unroll
            for (int j = 0; j < IT; j++) {
                const int k = s + lane + 32 * j;
                const bool in = k < e;
                f[j] = in ? a.flags[k] : (unsigned char)F_REMOVED;
                p[j] = in ? a.pos[k] : make_double2(0.0, 0.0);
            }
            // arrivals: lane j keeps the j-th one (from the records, or by walking the bag)
            int ax = -1, n_arr = 0;
            unsigned long long akey = ~0ull;
            if (n_rec <= kArrSlots) {
                n_arr = n_rec;
                if (lane < n_rec) {
                    const ArrRec r = a.arr_rec[(size_t)L * kArrSlots + lane];
                    ax = r.idx; akey = r.key; p[IT] = r.pos;
                }
            } else {
                for (int x = a.arr_head[L]; x >= 0 && n_arr <= 32; x = a.arr_next[x], n_arr++)
                    if (n_arr == lane) ax = x;
                if (ax >= 0) { p[IT] = a.pos[ax]; akey = entry_key(a, ax); }
            }
            if (n_arr > 32) {  // warp-uniform
                if (lane == 0) a.worklist[atomicAdd(a.work_count, 1)] = L;
                n_arr = -1;
            }
            if (n_arr >= 0) {
            reinterpret_cast<int4 *>(start)[2 * lane] = make_int4(0, 0, 0, 0);
            reinterpret_cast<int4 *>(start)[2 * lane + 1] = make_int4(0, 0, 0, 0);
            for (int o = lane; o < cap; o += 32) best[o] = 0xffffffffu;
            // order of the arrivals among themselves
            unsigned aord = 0;
            for (int t = 0; t < n_arr; t++) aord += __shfl_sync(0xffffffffu, akey, t) < akey ? 1u : 0u;
            if (ax >= 0) arr_idx[aord] = ax;
            __syncwarp();
            const int lrow = L / a.g.nc, gx = L - lrow * a.g.nc;
            const double x0 = gx * a.g.cellw, y0 = (a.g.row0 + lrow) * a.g.cellw;
            int br[IT + 1];  // bin << 16 | rank inside the bin
#pragma unroll
            for (int j = 0; j < IT; j++) {
                br[j] = -1;
                if (!(f[j] & F_REMOVED)) {
                    const int b = sub_bin16(a, p[j], x0, y0);
                    br[j] = (b << 16) | atomicAdd(&start[b], 1);
                } else if (f[j] & F_EMIGRANT) {
                    a.partner[s + lane + 32 * j] = -1;  // checked on the strip it moved to
                }
            }
            br[IT] = -1;
            if (ax >= 0) {
                const int b = sub_bin16(a, p[IT], x0, y0);
                br[IT] = (b << 16) | atomicAdd(&start[b], 1);
            }
            __syncwarp();
            warp_scan_bins(start, lane);
#pragma unroll
            for (int j = 0; j <= IT; j++) {
                if (br[j] >= 0) {
                    const int o = start[br[j] >> 16] + (br[j] & 0xffff);
                    br[j] = (br[j] & ~0xffff) | o;  // from here on: bin << 16 | slot in bin order
                    spos[o] = p[j];
                    sord[o] = j < IT ? (unsigned)(lane + 32 * j) : (0x80000000u | aord);
                }
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j <= IT; j++) {
                if (br[j] < 0) continue;
                const int b = br[j] >> 16;
                const int o = br[j] & 0xffff;
                const unsigned my = j < IT ? (unsigned)(lane + 32 * j) : (0x80000000u | aord);
                const int by = b >> 4, bx = b & 15;
                // rest of the own bin and the next bin of the row
                int q0 = o + 1, q1 = start[(by << 4) + min(bx + 1, a.nb - 1) + 1];
                for (int r = 0; r < 2; r++) {
                    tested += q1 - q0;
                    for (int q = q0; q < q1; q++) {
                        if (is_close(p[j], spos[q], a.thr2)) {
                            atomicMin(&best[o], sord[q]);
                            atomicMin(&best[q], my);
                        }
                    }
                    if (by + 1 >= a.nb) break;
                    q0 = start[((by + 1) << 4) + max(bx - 1, 0)];
                    q1 = start[((by + 1) << 4) + min(bx + 1, a.nb - 1) + 1];
                }
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j <= IT; j++) {
                if (br[j] < 0) continue;
                const int o = br[j] & 0xffff;
                const unsigned w = best[o];
                const int idx = w == 0xffffffffu ? -1 : (w & 0x80000000u) ? arr_idx[w & 0x7fffffffu] : s + (int)w;
                a.partner[j < IT ? s + lane + 32 * j : ax] = idx;
            }
            __syncwarp();
            }
        }
    }
    if (a.collision_pairs) block_add_u64((unsigned long long)tested, a.collision_pairs);
}

__global__ void __launch_bounds__(kThreads)
k_collide_big(CollideArgs a, int2 *__restrict__ huge_items, int *huge_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int scratch[kThreads / 32];
    const int nbins = a.nb * a.nb;
    Staging sm = carve(smem_raw, nbins, a.cap);
    long long tested = 0;
    const int nwork = *a.work_count;
    for (int w = blockIdx.x; w < nwork; w += gridDim.x) {
        const int L = a.worklist[w];
        long long r = collide_cell<kThreads>(a, L, threadIdx.x, sm, scratch);
        if (r < 0) {  // fits no staging area: slices for k_collide_huge
            if (threadIdx.x == 0) {
                const int nsl = (a.cell_start[L + 1] - a.cell_start[L] + kThreads - 1) / kThreads;
                const int base = atomicAdd(huge_count, nsl + 1);
                for (int j = 0; j < nsl; j++) huge_items[base + j] = make_int2(L, j);
                huge_items[base + nsl] = make_int2(L, -1);
            }
            r = 0;
        }
        tested += r;
    }
    if (a.collision_pairs) block_add_u64((unsigned long long)tested, a.collision_pairs);
}

}  // namespace parsim
