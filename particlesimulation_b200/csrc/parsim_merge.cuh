// parsim_merge.cuh -- end-of-step compaction and re-ordering without a sort (north_star (a)+(e)).
//
// Only ~1 % of the particles change cell per step (|v| dt <= 0.01 cell, serial/parsim.cpp:227-228)
// and the alive array is already ordered by (cell, id), so the order for the next step is built
// by a per-cell MERGE instead of a counting sort:
//   k_resolve : per particle -- who died (any close partner), which close pairs count as a
//               collision (mutual-first rule, see k_collide_find), and a one-byte fate; survivors
//               that stay are counted per cell with ONE atomic per (warp, cell); the few movers
//               are pushed on their destination cell's bag.
//   scan      : new cell_start / tombstone_start from the two count arrays.
//   k_move    : one lane group per destination cell -- copies the stayers of the old segment in
//               order (coalesced reads and writes), drops the dead, merges the bag by id, and
//               rebuilds the cell's tombstone list.  No atomics, deterministic layout.
#pragma once
#include <cooperative_groups.h>
#include <limits.h>

#include "parsim_kernels.cuh"

namespace parsim {
namespace cg = cooperative_groups;

enum Fate : unsigned char {
    FATE_STAY = 0,        // alive, same cell
    FATE_DEAD_HERE = 1,   // died, lies in the cell it was binned in
    FATE_MOVE = 2,        // alive, lies in another cell of this strip (in the move bag)
    FATE_DEAD_MOVED = 3,  // died, lies in another cell of this strip (in the dead bag)
    FATE_EMIGRATED = 4,   // handed to another strip before the collision phase
    FATE_LEFT_ALIVE = 5,  // alive, lies in another strip (sent by the late exchange)
    FATE_LEFT_DEAD = 6,   // died here, lies in another strip (its tombstone is sent)
};

// entries [0, n_res) are residents (cell[i] valid), [n_res, n) guests from other strips.
// Persistent blocks walk tiles of kThreads * kResolveRows entries, so the two contended global
// counters cost one atomic per TILE (graveyard slots) and one per BLOCK (collision total)
// instead of one per warp -- same-address atomics serialise in L2 at about 1 ns each.
#ifndef PS_RESOLVE_ROWS
#define PS_RESOLVE_ROWS 4  // measured: 4 rows at 8 blocks per SM (32 registers) beat 8 rows at 4 blocks by 15-25 %
#endif
#ifndef PS_RESOLVE_MINB
#define PS_RESOLVE_MINB 8
#endif
constexpr int kResolveRows = PS_RESOLVE_ROWS;

__global__ void __launch_bounds__(kThreads, PS_RESOLVE_MINB)
k_resolve(Geom g, int n_res_val, const int *__restrict__ n_res_ptr, int n_guest, const int *__restrict__ partner, const int *__restrict__ cell,
          const int *__restrict__ newcell, const unsigned char *__restrict__ flags,
          const double2 *__restrict__ pos, const double2 *__restrict__ vel,
          const unsigned *__restrict__ id, unsigned char *__restrict__ fate, int *count,
          int *dead_new_count, int *mv_head, int *dm_head, int *__restrict__ mv_next,
          double2 *__restrict__ grave_pos, double2 *__restrict__ grave_vel,
          unsigned *__restrict__ grave_id, int *__restrict__ grave_cell, int grave_cap, Counters *ctr,
          int *__restrict__ late_list, int *late_count, int late_cap) {
    constexpr int kWarps = kThreads / 32;
    const int n_res = n_res_ptr ? *n_res_ptr : n_res_val, n = n_res + n_guest;
    __shared__ int warp_dying[kResolveRows][kWarps];
    __shared__ int tile_base;
    __shared__ unsigned warp_cnt[kWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned mine = 0;
    const int tile_len = kThreads * kResolveRows;
    for (long long t0 = (long long)blockIdx.x * tile_len; t0 < n; t0 += (long long)gridDim.x * tile_len) {
        unsigned char fts[kResolveRows];
        unsigned dying_mask[kResolveRows];
        // all loads of the tile are issued before the first one is needed, in two waves
        // (partner[partner[i]] depends on the first wave)
        int cc[kResolveRows], nn[kResolveRows], pp[kResolveRows], qq[kResolveRows];
        unsigned char ff[kResolveRows];
#pragma unroll
        for (int r = 0; r < kResolveRows; r++) {
            const long long ii = t0 + (long long)r * kThreads + threadIdx.x;
            const bool in = ii < n;
            const int i = (int)ii;
            cc[r] = in && i < n_res ? cell[i] : -1;
            nn[r] = in ? newcell[i] : 0;
            ff[r] = in ? flags[i] : (unsigned char)0;
            pp[r] = in ? partner[i] : -1;
        }
#pragma unroll
        for (int r = 0; r < kResolveRows; r++) qq[r] = pp[r] >= 0 ? partner[pp[r]] : -1;
#pragma unroll
        for (int r = 0; r < kResolveRows; r++) {
            const long long ii = t0 + (long long)r * kThreads + threadIdx.x;
            const int i = (int)ii;
            int c = -2 - lane;  // lanes past the end match nobody
            int ft = 255, nl = -1;
            if (ii < n) {
                c = i < n_res ? cc[r] : -40 - lane;  // guests have no bin-time cell
                nl = nn[r] - g.cell_base;
                const bool in_strip = nl >= 0 && nl < g.ncells;
                if (ff[r] & F_EMIGRANT) {
                    ft = FATE_EMIGRATED;
                } else {
                    const int p = pp[r];
                    const bool dead = p >= 0;
                    if (dead && qq[r] == i && i < p) mine++;  // mutual-first pair, counted once
                    if (!in_strip) {  // multi-strip only: packed for the late exchange from this short list
                        ft = dead ? FATE_LEFT_DEAD : FATE_LEFT_ALIVE;
                        const int sl = atomicAdd(late_count, 1);
                        if (sl < late_cap) late_list[sl] = i;
                        else ctr->error = DE_MIGRANT_OVERFLOW;
                    }
                    else if (nl == c) ft = dead ? FATE_DEAD_HERE : FATE_STAY;
                    else ft = dead ? FATE_DEAD_MOVED : FATE_MOVE;
                }
                fate[i] = (unsigned char)ft;
            }
            // per-cell counts: the array is cell-ordered, so a warp touches 1-3 cells
            const unsigned grp = __match_any_sync(0xffffffffu, c);
            const unsigned stay = __ballot_sync(0xffffffffu, ft == FATE_STAY);
            const unsigned dhere = __ballot_sync(0xffffffffu, ft == FATE_DEAD_HERE);
            if (c >= 0 && lane == __ffs(grp) - 1) {
                int ns = __popc(grp & stay), nd = __popc(grp & dhere);
                if (ns) atomicAdd(&count[c], ns);
                if (nd) atomicAdd(&dead_new_count[c], nd);
            }
            if (ft == FATE_MOVE) {
                atomicAdd(&count[nl], 1);
                mv_next[i] = atomicExch(&mv_head[nl], i);
            } else if (ft == FATE_DEAD_MOVED) {
                atomicAdd(&dead_new_count[nl], 1);
                mv_next[i] = atomicExch(&dm_head[nl], i);
            }
            const bool dying = ft == FATE_DEAD_HERE || ft == FATE_DEAD_MOVED || ft == FATE_LEFT_DEAD;
            const unsigned dm = __ballot_sync(0xffffffffu, dying);
            fts[r] = (unsigned char)ft;
            dying_mask[r] = dm;
            if (lane == 0) warp_dying[r][warp] = __popc(dm);
        }
        __syncthreads();
        // graveyard (frozen state of the dead, for parsim_result / dumps): one reservation per tile
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int r = 0; r < kResolveRows; r++)
                for (int w = 0; w < kWarps; w++) { int v = warp_dying[r][w]; warp_dying[r][w] = tot; tot += v; }
            tile_base = tot ? atomicAdd(&ctr->grave_count, tot) : 0;
            if (tile_base + tot > grave_cap) ctr->error = DE_CAPACITY;  // more deaths than this strip has slots
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < kResolveRows; r++) {
            const int ft = fts[r];
            if (ft == FATE_DEAD_HERE || ft == FATE_DEAD_MOVED || ft == FATE_LEFT_DEAD) {
                const int i = (int)(t0 + (long long)r * kThreads + threadIdx.x);
                const int slot = tile_base + warp_dying[r][warp] + __popc(dying_mask[r] & ((1u << lane) - 1u));
                if (slot >= grave_cap) continue;  // reported through ctr->error above
                grave_pos[slot] = pos[i];
                grave_vel[slot] = vel[i];
                grave_id[slot] = id[i];
                grave_cell[slot] = newcell[i];
            }
        }
        __syncthreads();  // warp_dying / tile_base are reused by the next tile
    }
    unsigned w = __reduce_add_sync(0xffffffffu, mine);
    if (lane == 0) warp_cnt[warp] = w;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned tot = 0;
#pragma unroll
        for (int k = 0; k < kWarps; k++) tot += warp_cnt[k];
        if (tot) atomicAdd(&ctr->collisions, (unsigned long long)tot);
    }
}

struct MoveArgs {
    const int *cs_old, *cs_new;          // cell_start before / after
    const int *ds_old, *ds_new;          // tombstone_start before / after (ds_new == nullptr: no dead yet)
    const unsigned char *fate;
    const double2 *pos_in, *vel_in;      // state of this step, old order (+ guests / immigrants behind)
    const double *m_in;
    const unsigned *id_in;
    int *mv_head, *dm_head, *arr_head;   // bags; reset to -1 here for the next step
    int *arr_cnt;                        // arrival-record counters (nullptr: unused); reset to 0 here
    int *count;                          // per-cell survivor counts; reset to 0 here
    const int *mv_next;
    const unsigned *dead_old;
    double2 *pos_out, *vel_out;
    double *m_out;
    unsigned *id_out;
    int *cell_out;
    unsigned *dead_new;
    int *big_list, *big_count;           // cells with more than big_min old entries: left to k_move_big
    int big_min;
    double2 *com_xy;                     // COM variants only: next step's centres of mass, already
    double *com_m;                       // offset past the ghost row (index = local cell)
};

// BIG: leave cells (or tombstone lists) above a.big_min entries to k_move_big.  The test costs the
// ordinary case about 5 % (this kernel is a chain of dependent loads per cell), so the host asks
// for it only when the cells are huge on average.
//
// COM: the lane group also leaves the cell's centre of mass for the NEXT step (k_com's job, saved
// with its 24 B per particle of DRAM reads): it reads its own freshly written list back from L2
// and sums it in exactly k_com<LANES>'s order -- lane l takes entries l, l + LANES, ... with the
// same fma chain, then the same xor tree and divisions -- so the result is k_com's bit for bit.
// The host uses it only when LANES is the context's com_lanes (the lane count fixes the order).
template <int LANES, bool BIG, bool COM>
__device__ __forceinline__ void move_cell(const Geom &g, const MoveArgs &a, int c) {
    auto tile = cg::tiled_partition<LANES>(cg::this_thread_block());
    const int lane = tile.thread_rank();
    const unsigned lt = (1u << lane) - 1u;
    const int s = a.cs_old[c], e = a.cs_old[c + 1];
    const int ns = a.cs_new[c];
    if (BIG) {
        int nd = 0;
        if (a.ds_new) nd = a.ds_old[c + 1] - a.ds_old[c];
        if (max(e - s, nd) > a.big_min) {  // would keep this one lane group busy for the whole kernel
            if (lane == 0) a.big_list[atomicAdd(a.big_count, 1)] = c;
            return;
        }
    }

    // the first LANES arrivals live in registers (lane j holds arrival j); more are walked
    int n_arr = 0, my_arr = -1, rest = -1;
    unsigned my_aid = 0;
    for (int x = a.mv_head[c]; x >= 0; x = a.mv_next[x]) {
        if (n_arr == LANES) { rest = x; break; }
        if (lane == n_arr) { my_arr = x; my_aid = a.id_in[x]; }
        n_arr++;
    }

    // tombstones: old entries first
    int doff = 0;
    if (a.ds_new) {
        const int ds = a.ds_old[c], de = a.ds_old[c + 1];
        doff = a.ds_new[c];
        for (int t = ds + lane; t < de; t += LANES) a.dead_new[doff + (t - ds)] = a.dead_old[t];
        doff += de - ds;
    }

    // stayers keep their order; each adds the number of arrivals with a smaller id
    int run = 0;
    for (int base = s; base < e; base += LANES) {
        const int k = base + lane;
        const int ft = k < e ? a.fate[k] : 255;
        const unsigned me = k < e ? a.id_in[k] : 0u;
        const unsigned smask = tile.ballot(ft == FATE_STAY);
        int before = 0;  // arrivals with a smaller id (shuffles stay outside divergent code)
        for (int j = 0; j < n_arr; j++) before += tile.shfl(my_aid, j) < me ? 1 : 0;
        if (ft == FATE_STAY) {
            // the pointers of MoveArgs may alias as far as the compiler knows, so loads and stores
            // issue in program order: the whole payload is in flight before the first store
            const double2 pp = a.pos_in[k], vv = a.vel_in[k];
            const double mm = a.m_in[k];
            int r = run + __popc(smask & lt) + before;
            for (int x = rest; x >= 0; x = a.mv_next[x]) r += a.id_in[x] < me ? 1 : 0;
            const int o = ns + r;
            a.pos_out[o] = pp;
            a.vel_out[o] = vv;
            a.m_out[o] = mm;
            a.id_out[o] = me;
            a.cell_out[o] = c;
        }
        run += __popc(smask);
        if (a.ds_new) {
            const unsigned dmask = tile.ballot(ft == FATE_DEAD_HERE);
            if (ft == FATE_DEAD_HERE) a.dead_new[doff + __popc(dmask & lt)] = me;
            doff += __popc(dmask);
        }
    }

    // arrivals: position = stayers with a smaller id + arrivals with a smaller id
    if (rest < 0) {
        // all of them are in registers (lane q holds arrival q): no second walk of the bag, and
        // the lanes move their arrivals together (one round of loads, then the stores)
        int myo = -1;
        for (int q = 0; q < n_arr; q++) {
            const unsigned aid = tile.shfl(my_aid, q);
            int r = __popc(tile.ballot(lane < n_arr && my_aid < aid));
            for (int base = s; base < e; base += LANES) {
                const int k = base + lane;
                const bool less = k < e && a.fate[k] == FATE_STAY && a.id_in[k] < aid;
                r += __popc(tile.ballot(less));
            }
            if (lane == q) myo = ns + r;
        }
        if (myo >= 0) {
            const double2 pp = a.pos_in[my_arr], vv = a.vel_in[my_arr];
            const double mm = a.m_in[my_arr];
            a.pos_out[myo] = pp;
            a.vel_out[myo] = vv;
            a.m_out[myo] = mm;
            a.id_out[myo] = my_aid;
            a.cell_out[myo] = c;
        }
    } else {
        for (int x = a.mv_head[c]; x >= 0; x = a.mv_next[x]) {
            const unsigned aid = a.id_in[x];
            int r = 0;
            for (int base = s; base < e; base += LANES) {
                const int k = base + lane;
                const bool less = k < e && a.fate[k] == FATE_STAY && a.id_in[k] < aid;
                r += __popc(tile.ballot(less));
            }
            for (int y = a.mv_head[c]; y >= 0; y = a.mv_next[y]) r += a.id_in[y] < aid ? 1 : 0;
            if (lane == 0) {
                const int o = ns + r;
                a.pos_out[o] = a.pos_in[x];
                a.vel_out[o] = a.vel_in[x];
                a.m_out[o] = a.m_in[x];
                a.id_out[o] = aid;
                a.cell_out[o] = c;
            }
        }
    }
    // tombstones of particles that died this step and came to rest in this cell from elsewhere
    if (a.ds_new && lane == 0)
        for (int x = a.dm_head[c]; x >= 0; x = a.mv_next[x]) a.dead_new[doff++] = a.id_in[x];
    tile.sync();
    if (lane == 0) {
        a.mv_head[c] = -1;
        a.dm_head[c] = -1;
        a.arr_head[c] = -1;
        if (a.arr_cnt) a.arr_cnt[c] = 0;
        a.count[c] = 0;
    }
    if (COM) {
        // tile.sync() above ordered the lanes' stores before these loads; .cg reads L2, where the
        // stores went (lines shared with a neighbouring cell may be stale in L1 for ITS part only)
        const int ne = a.cs_new[c + 1];
        double M = 0, Mx = 0, My = 0;
        for (int k = ns + lane; k < ne; k += LANES) {
            const double mv = __ldcg(a.m_out + k);
            const double2 p = __ldcg(a.pos_out + k);
            M += mv;
            Mx = fma(mv, p.x, Mx);
            My = fma(mv, p.y, My);
        }
#pragma unroll
        for (int o = LANES / 2; o > 0; o >>= 1) {
            M += tile.shfl_xor(M, o);
            Mx += tile.shfl_xor(Mx, o);
            My += tile.shfl_xor(My, o);
        }
        // lanes 0 and 1 divide one component each (an FP64 division is ~40 instructions)
        double val = 0.0;
        if (lane < 2 && M > 0) val = (lane ? My : Mx) / M;
        const double other = tile.shfl_down(val, 1);
        if (lane == 0) {
            a.com_xy[c] = make_double2(val, other);
            a.com_m[c] = M;
        }
    }
}

template <int LANES, bool BIG, bool COM>
__global__ void __launch_bounds__(kThreads, 8) k_move(Geom g, MoveArgs a) {
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) / LANES;
    if (c >= g.ncells) return;  // whole tiles leave together
    move_cell<LANES, BIG, COM>(g, a, c);
}

__device__ __noinline__ void move_cell_lanes32(const Geom &g, const MoveArgs &a, int c) { move_cell<32, false, false>(g, a, c); }

// Huge cells (tens of thousands of entries, BASELINE config 5): one BLOCK per cell.  The arrivals
// (<= kMoveBigArr, else the block falls back to a single lane group's generic loop) are sorted by
// id in shared memory; a stayer's slot is (stayers before it) + (arrivals with a smaller id, by
// binary search); the histogram of those search results, prefix-summed, gives every arrival its
// slot without looking at the stayers again.
constexpr int kMoveBigArr = 2048;

__global__ void __launch_bounds__(kThreads) k_move_big(Geom g, MoveArgs a) {
    __shared__ unsigned a_id[kMoveBigArr], s_id[kMoveBigArr];
    __shared__ int a_ix[kMoveBigArr], s_ix[kMoveBigArr];
    __shared__ int hist[kMoveBigArr + 1];
    __shared__ int wsum[kThreads / 32], wdead[kThreads / 32];
    __shared__ int sh_n_arr, sh_run, sh_doff;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int nbig = *a.big_count;
    for (int w = blockIdx.x; w < nbig; w += gridDim.x) {
        const int c = a.big_list[w];
        const int s = a.cs_old[c], e = a.cs_old[c + 1];
        const int ns = a.cs_new[c];
        if (t == 0) {
            int n = 0;
            for (int x = a.mv_head[c]; x >= 0; x = a.mv_next[x], n++)
                if (n < kMoveBigArr) a_ix[n] = x;
            sh_n_arr = n;
        }
        __syncthreads();
        const int n_arr = sh_n_arr;
        if (n_arr > kMoveBigArr) {  // block-uniform; never seen in practice (1 % of a cell moves per step)
            // the generic loop needs no cooperation beyond one lane group: run it on the first warp
            // with the big-cell test switched off
            if (warp == 0) move_cell_lanes32(g, a, c);
            __syncthreads();
            continue;
        }
        for (int j = t; j < n_arr; j += kThreads) a_id[j] = a.id_in[a_ix[j]];
        for (int j = t; j <= n_arr; j += kThreads) hist[j] = 0;
        __syncthreads();
        for (int j = t; j < n_arr; j += kThreads) {  // rank sort by id (ids are distinct)
            const unsigned me = a_id[j];
            int r = 0;
            for (int q = 0; q < n_arr; q++) r += a_id[q] < me ? 1 : 0;
            s_id[r] = me;
            s_ix[r] = a_ix[j];
        }
        // tombstones: old entries first
        int doff = 0;
        if (a.ds_new) {
            const int ds = a.ds_old[c], de = a.ds_old[c + 1];
            doff = a.ds_new[c];
            for (int q = ds + t; q < de; q += kThreads) a.dead_new[doff + (q - ds)] = a.dead_old[q];
            doff += de - ds;
        }
        if (t == 0) { sh_run = 0; sh_doff = doff; }
        __syncthreads();
        for (int base = s; base < e; base += kThreads) {
            const int k = base + t;
            const int ft = k < e ? a.fate[k] : 255;
            const unsigned me = k < e ? a.id_in[k] : 0u;
            const bool stay = ft == FATE_STAY, dead = ft == FATE_DEAD_HERE;
            const unsigned smask = __ballot_sync(0xffffffffu, stay), dmask = __ballot_sync(0xffffffffu, dead);
            if (lane == 0) { wsum[warp] = __popc(smask); wdead[warp] = __popc(dmask); }
            __syncthreads();
            int sbefore = 0, dbefore = 0, stot = 0, dtot = 0;
#pragma unroll
            for (int q = 0; q < kThreads / 32; q++) {
                sbefore += q < warp ? wsum[q] : 0;
                dbefore += q < warp ? wdead[q] : 0;
                stot += wsum[q];
                dtot += wdead[q];
            }
            const int run = sh_run, dbase = sh_doff;
            if (stay) {
                int lo = 0, hi = n_arr;  // arrivals with a smaller id
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (s_id[mid] < me) lo = mid + 1; else hi = mid;
                }
                atomicAdd(&hist[lo], 1);
                const int o = ns + run + sbefore + __popc(smask & lt) + lo;
                a.pos_out[o] = a.pos_in[k];
                a.vel_out[o] = a.vel_in[k];
                a.m_out[o] = a.m_in[k];
                a.id_out[o] = me;
                a.cell_out[o] = c;
            }
            if (dead && a.ds_new) a.dead_new[dbase + dbefore + __popc(dmask & lt)] = me;
            __syncthreads();
            if (t == 0) { sh_run = run + stot; sh_doff = dbase + dtot; }
            __syncthreads();
        }
        // arrival j (id order): j arrivals and hist[0..j] stayers lie before it
        if (t == 0) {
            int acc = 0;
            for (int j = 0; j <= n_arr; j++) { acc += hist[j]; hist[j] = acc; }
        }
        __syncthreads();
        for (int j = t; j < n_arr; j += kThreads) {
            const int x = s_ix[j];
            const int o = ns + j + hist[j];
            a.pos_out[o] = a.pos_in[x];
            a.vel_out[o] = a.vel_in[x];
            a.m_out[o] = a.m_in[x];
            a.id_out[o] = s_id[j];
            a.cell_out[o] = c;
        }
        if (t == 0) {
            int d = sh_doff;
            if (a.ds_new)
                for (int x = a.dm_head[c]; x >= 0; x = a.mv_next[x]) a.dead_new[d++] = a.id_in[x];
            a.mv_head[c] = -1;
            a.dm_head[c] = -1;
            a.arr_head[c] = -1;
            if (a.arr_cnt) a.arr_cnt[c] = 0;
            a.count[c] = 0;
        }
        __syncthreads();
    }
}

}  // namespace parsim
