// parsim_sort.cuh -- binning of unordered particles into the cell-ordered SoA layout
// (north_star (a); the reference's updateCOM loop, serial/parsim.cpp:265-285, appends particle i
// to cellParticles[cell(i)] for i ascending, i.e. every list is ordered by particle index).
//
// A stable LSD radix sort of (cell key, source slot) pairs, 7-8 bits per pass:
//   k_sort_hist    : per-tile digit histogram                      (reads 4 B/p)
//   scan           : digit-major exclusive scan of the tile histograms (parsim_kernels.cuh)
//   k_sort_scatter : stable in-tile ranking with warp match/popc, tile staged in shared memory in
//                    digit order, then written out in runs          (reads 8 B/p, writes 8 B/p)
//   k_sort_gather  : moves the 44-byte state once through the final permutation and derives
//                    cell_start from the sorted keys (no histogram, no atomics)
// The sort is stable and the sources are visited in slot order, so when the slots are in id
// order (parsim_upload / parsim_init_device) each cell comes out ordered by id -- exactly the
// reference's bin-time list.  When they are not (explicit-id uploads in arbitrary order), the id
// digits are sorted first (LSD), which costs more passes but gives the same layout.
#pragma once
#include "parsim_kernels.cuh"

namespace parsim {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 entries per block
constexpr int kSortMaxBins = 256;

__device__ __forceinline__ unsigned sort_digit(unsigned key, int shift, unsigned mask) {
    return (key >> shift) & mask;
}

// hist layout: hist[d * ntiles + tile]
__global__ void __launch_bounds__(kSortThreads)
k_sort_hist(int n, const unsigned *__restrict__ keys, int shift, unsigned mask, int ntiles,
            int *__restrict__ hist) {
    __shared__ int h[kSortMaxBins];
    for (int d = threadIdx.x; d <= (int)mask; d += kSortThreads) h[d] = 0;
    __syncthreads();
    const long long base = (long long)blockIdx.x * kSortTile;
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        long long i = base + j * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[sort_digit(keys[i], shift, mask)], 1);
    }
    __syncthreads();
    for (int d = threadIdx.x; d <= (int)mask; d += kSortThreads) hist[(size_t)d * ntiles + blockIdx.x] = h[d];
}

// src_in == nullptr: the source of entry i is i (first pass).
#ifndef PS_SCATTER_MINB
#define PS_SCATTER_MINB 3  // 85 registers; measured best (1: 4.4 ms, 3: 2.7, 4: 2.9, 5: 3.7 for three passes over 1e8 keys)
#endif
__global__ void __launch_bounds__(kSortThreads, PS_SCATTER_MINB)
k_sort_scatter(int n, const unsigned *__restrict__ keys_in, const int *__restrict__ src_in, int shift,
               unsigned mask, int ntiles, const int *__restrict__ hist_scanned,
               unsigned *__restrict__ keys_out, int *__restrict__ src_out) {
    __shared__ int whist[kSortWarps][kSortMaxBins];  // per-warp digit counts, then warp bases
    __shared__ int bstart[kSortMaxBins];             // start of each digit in the tile-sorted order
    __shared__ int gbase[kSortMaxBins];              // global start of this tile's run of each digit
    __shared__ int wsum[kSortWarps];
    __shared__ unsigned skey[kSortTile];
    __shared__ int ssrc[kSortTile];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int nb = (int)mask + 1;
    for (int d = threadIdx.x; d < kSortWarps * kSortMaxBins; d += kSortThreads) (&whist[0][0])[d] = 0;
    __syncthreads();

    // warp w owns the contiguous slice [base + w*512, base + (w+1)*512); item j of lane l is
    // entry j*32 + l of the slice, so (j, lane) order is index order (stability)
    const long long wbase = (long long)blockIdx.x * kSortTile + (long long)warp * (32 * kSortItems);
    unsigned key[kSortItems];
    int src[kSortItems];
    int rank[kSortItems];
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        const long long i = wbase + j * 32 + lane;
        const bool ok = i < n;
        key[j] = ok ? keys_in[i] : 0u;
        src[j] = ok ? (src_in ? src_in[i] : (int)i) : -1;
    }
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        const bool ok = src[j] >= 0;
        const int d = ok ? (int)sort_digit(key[j], shift, mask) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        int cnt = 0;
        if (ok) cnt = whist[warp][d];
        __syncwarp();
        rank[j] = cnt + __popc(peers & lt);
        if (ok && (peers & lt) == 0) whist[warp][d] = cnt + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // per digit: exclusive prefix over the warps (warp order = index order), tile total
    int tot = 0;
    if (threadIdx.x < nb) {
        const int d = threadIdx.x;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            int v = whist[w][d];
            whist[w][d] = tot;
            tot += v;
        }
        gbase[d] = hist_scanned[(size_t)d * ntiles + blockIdx.x];
    }
    // exclusive scan of the tile totals over the digits (nb <= 256 = one value per thread)
    int inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    int carry = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; w++) carry += (w < warp) ? wsum[w] : 0;
    if (threadIdx.x < nb) bstart[threadIdx.x] = carry + inc - tot;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        if (src[j] >= 0) {
            const int d = (int)sort_digit(key[j], shift, mask);
            const int p = bstart[d] + whist[warp][d] + rank[j];
            skey[p] = key[j];
            ssrc[p] = src[j];
        }
    }
    __syncthreads();
    const long long tile_first = (long long)blockIdx.x * kSortTile;
    const int count = (int)min((long long)kSortTile, (long long)n - tile_first);
    for (int p = threadIdx.x; p < count; p += kSortThreads) {
        const unsigned k = skey[p];
        const int d = (int)sort_digit(k, shift, mask);
        const int o = gbase[d] + (p - bstart[d]);
        keys_out[o] = k;
        src_out[o] = ssrc[p];
    }
}

// keys for an id pass: the id of the source of entry i (src_in == nullptr: source = i)
__global__ void k_sort_id_keys(int n, const int *__restrict__ src_in, const InitRec *__restrict__ rec,
                               unsigned *__restrict__ keys) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = rec[src_in ? src_in[i] : i].id;
}
// keys for the first cell pass after id passes: the cell of the source of entry i
__global__ void k_sort_cell_keys(int n, const int *__restrict__ src_in, const unsigned *__restrict__ cellkey,
                                 unsigned *__restrict__ keys) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = cellkey[src_in[i]];
}

// are the record ids ascending in slot order?  (*flag is set to 1 when they are not)
__global__ void k_rec_ids_ascending(int n, const InitRec *__restrict__ rec, int *flag) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 < n && rec[i].id >= rec[i + 1].id) *flag = 1;
}

// cell key of every slot (the reference's binning arithmetic, :268-272)
__global__ void k_sort_cell_of(Geom g, int n, const InitRec *__restrict__ rec, unsigned *__restrict__ key) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) key[i] = (unsigned)(global_cell(rec[i].pos, g) - g.cell_base);
}

// final permutation: one 64-byte read per particle, SoA written in (cell, id) order;
// cell_start from the sorted keys
__global__ void __launch_bounds__(kThreads)
k_sort_gather(int n, int ncells, const unsigned *__restrict__ keys, const int *__restrict__ src,
              const InitRec *__restrict__ rec, double2 *__restrict__ pos_out, double2 *__restrict__ vel_out,
              double *__restrict__ m_out, unsigned *__restrict__ id_out, int *__restrict__ cell_out,
              int *__restrict__ cell_start, int *__restrict__ where = nullptr) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    if (where) where[src[p]] = p;  // slot -> ordered position (late velocities of a strip upload)
    const int c = (int)keys[p];
    const double2 *q = reinterpret_cast<const double2 *>(rec + src[p]);
    const double2 a = __ldg(q), v = __ldg(q + 1), mi = __ldg(q + 2);
    pos_out[p] = a;
    vel_out[p] = v;
    m_out[p] = mi.x;
    id_out[p] = (unsigned)__double2loint(mi.y);
    cell_out[p] = c;
    const int prev = p > 0 ? (int)keys[p - 1] : -1;
    for (int k = prev + 1; k <= c; k++) cell_start[k] = p;
    if (p == n - 1)
        for (int k = c + 1; k <= ncells; k++) cell_start[k] = n;
}

// ---------------------------------------------------------------------------------------
// Upload pieces sorted under the copy's shadow (parsim_step.inl sort_piece / merge_pieces): every
// piece of the upload is already in (cell, id) order inside itself, pieces hold increasing id
// ranges, so the (cell, id) order of the whole is, per cell, piece 0's segment, then piece 1's, ...
// cs_pieces[p * (ncells + 1) + c] = start of cell c inside piece p (relative to piece_first[p]).
__global__ void k_piece_counts(int ncells, int npieces, const int *__restrict__ cs_pieces, int *__restrict__ count) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    int tot = 0;
    for (int p = 0; p < npieces; p++) {
        const int *cs = cs_pieces + (size_t)p * (ncells + 1);
        tot += cs[c + 1] - cs[c];
    }
    count[c] = tot;
}

// MAXP > 0: at most MAXP pieces; all segment bounds of the cell are loaded before the first copy
// (2 * npieces independent loads in flight instead of a dependent pair per piece).
template <int LANES, int MAXP>
__global__ void __launch_bounds__(kThreads, MAXP > 0 ? 5 : 8)
k_merge_pieces(int ncells, int npieces, const int *__restrict__ piece_first, const int *__restrict__ cs_pieces,
               const int *__restrict__ cs_new, const double2 *__restrict__ pos_in, const double2 *__restrict__ vel_in,
               const double *__restrict__ m_in, const unsigned *__restrict__ id_in, double2 *__restrict__ pos_out,
               double2 *__restrict__ vel_out, double *__restrict__ m_out, unsigned *__restrict__ id_out,
               int *__restrict__ cell_out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = t / LANES, lane = t % LANES;
    if (c >= ncells) return;
    int dst = cs_new[c];
    int ss[MAXP > 0 ? MAXP : 1], ee[MAXP > 0 ? MAXP : 1];
    if (MAXP > 0) {
#pragma unroll
        for (int p = 0; p < MAXP; p++) {
            const int *cs = cs_pieces + (size_t)p * (ncells + 1);
            ss[p] = p < npieces ? cs[c] : 0;
            ee[p] = p < npieces ? cs[c + 1] : 0;
        }
    }
#pragma unroll
    for (int p = 0; p < (MAXP > 0 ? MAXP : npieces); p++) {
        if (MAXP > 0 && p >= npieces) break;
        int s, e;
        if (MAXP > 0) {
            s = ss[p]; e = ee[p];
        } else {
            const int *cs = cs_pieces + (size_t)p * (ncells + 1);
            s = cs[c]; e = cs[c + 1];
        }
        if (e > s) {
            const int base = piece_first[p];
            for (int k = s + lane; k < e; k += LANES) {
                const int i = base + k, o = dst + (k - s);
                const double2 pp = pos_in[i], vv = vel_in[i];
                const double mm = m_in[i];
                const unsigned id = id_in[i];
                pos_out[o] = pp;
                vel_out[o] = vv;
                m_out[o] = mm;
                id_out[o] = id;
                cell_out[o] = c;
            }
        }
        dst += e - s;
    }
}

}  // namespace parsim
