// parsim_rank.cuh -- host-side state of ONE row strip on ONE GPU ("rank"), and the per-step
// driver that launches the kernels of parsim_kernels.cuh in the reference's phase order
// (ParticleSimulation::simulate, serial/parsim.cpp:432-441).
#pragma once
#include <cuda_runtime.h>

#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/parsim_b200.h"
#include "parsim_kernels.cuh"
#include "parsim_merge.cuh"
#include "parsim_sort.cuh"
#include "parsim_collide.cuh"
#include "parsim_nccl.hpp"

namespace parsim {

enum KernelId {
    K_INGEST = 0, K_BIN_KEYS, K_SCAN, K_SORT_PASS, K_SORT_GATHER, K_COM, K_GHOST, K_FORCE,
    K_CLASSIFY, K_COLLIDE, K_RESOLVE, K_TOMBSTONES, K_MIGRATE, K_INIT, K_MOVE, K_COLLIDE_BIG, K_COUNT_
};
static_assert(K_COUNT_ <= PARSIM_NKERNELS, "grow PARSIM_NKERNELS");

struct TimedLaunch {
    cudaEvent_t a, b;
    int k;
};

struct Rank;

// In-process transport of an oversubscribed / peer-copy group (parsim_create with more strips than
// GPUs, or PARSIM_TRANSPORT=local): the strips of ONE process exchange by device-to-device copies
// out of each other's send buffers instead of ncclSend/ncclRecv -- the reference's own way of
// testing its distributed variant on one machine (mpi/run_tests.sh:9-17: oversubscribed local
// ranks).  Only the transport differs; pack / unpack kernels and the order rules are the NCCL
// path's.  One host thread per strip (parsim_capi.cu for_ranks) meets the others at barrier();
// a strip that failed breaks the barrier so that nobody waits for it for ever.
struct LocalBus {
    std::vector<Rank *> ranks;
    std::mutex mu;
    std::condition_variable cv;
    int waiting = 0;
    unsigned long long generation = 0;
    bool failed = false;

    int barrier() {  // 0 = everybody arrived, 1 = some strip failed
        std::unique_lock<std::mutex> lk(mu);
        if (failed) return 1;
        const unsigned long long gen = generation;
        if (++waiting == (int)ranks.size()) {
            waiting = 0;
            generation++;
            cv.notify_all();
            return 0;
        }
        cv.wait(lk, [&] { return generation != gen || failed; });
        return failed && generation == gen ? 1 : 0;
    }
    void abort() {
        std::lock_guard<std::mutex> lk(mu);
        failed = true;
        cv.notify_all();
    }
    void rearm() {  // before the next API call (a failed call leaves the context unusable for steps anyway)
        std::lock_guard<std::mutex> lk(mu);
        waiting = 0;
    }
};

constexpr int kCounterRing = 2;  // host copies of the device counters of the last two steps (deferred mode)

struct Rank {
    // problem
    double side = 0;
    long nc = 0;
    long long n_part = 0;
    int rank = 0, nranks = 1, device = 0;
    Geom g{};
    double thr2 = 0;    // collision threshold on the squared distance (see k_collide_find)
    int com_lanes = 32; // lanes per cell in k_com, from the GLOBAL particles-per-cell
    int force_variant = 0; // k_force_integrate instantiation (PARSIM_FORCE_VARIANT; 0: by occupancy)
    int sparse_variant = 5; // the instantiation for sparse cells (PARSIM_SPARSE_VARIANT: 1 or 5)

    // particle storage: alive residents [0, n_alive), guests after them until the rebin
    int cap = 0;
    int n_alive = 0;
    int n_guest = 0;
    bool sorted = false;   // cell order valid for the coming step
    bool count_pending = false;  // several strips: n_alive of the coming step is still on its way (ev_ring[0], h_ring[0])
    double2 *acc0 = nullptr;     // accelerations of the coming step, taken during the upload (late velocities)
    double *late_vx = nullptr, *late_vy = nullptr;  // the velocity arrays of such an upload, slot order
    bool halo_valid = false;       // the ghost rows of the coming step are in place (collective strip upload)
    int *up_slot = nullptr;        // strip upload with late velocities: slot of every uploaded particle (-1: not mine)
    size_t up_slot_cap = 0, late_cap = 0;
    long long late_count = 0;      // entries of late_vx / late_vy / up_slot the pending record patch covers (0: slot order, n_part)
    bool rec_vel_pending = false;  // the records' velocities are still in late_vx / late_vy (written on demand: reset, full sort)
    bool opt_vel_late = true;    // PARSIM_VEL_LATE=0: velocities travel with every piece (A/B)
    bool com_valid = false;  // the merge of the last step already left this step's centres of mass (k_move<.., COM>)
    bool fuse_com = true;    // PARSIM_FUSE_COM=0: always k_com (A/B)
    int next_stage = 0;    // debug stepping
    double2 *pos_cur = nullptr, *pos_alt = nullptr;  // cur: pre-move (sorted); alt: post-move
    double2 *vel_cur = nullptr, *vel_alt = nullptr;
    double *m_cur = nullptr, *m_alt = nullptr;
    unsigned *id_cur = nullptr, *id_alt = nullptr;
    int *cell_cur = nullptr, *cell_alt = nullptr;
    int *cs_cur = nullptr, *cs_alt = nullptr;  // cell_start[ncells+1]
    double2 *pos_latest = nullptr;             // where the newest positions are (dumps)
    float2 *posf = nullptr;                    // FP32 copy of the post-move positions (collision pre-filter)
    float thrf = 0;                            // pre-filter threshold on the FP32 squared distance
    // cell-centric collision search (parsim_collide.cuh); off: per-particle all-pairs k_collide_find
    bool collide_grid = false;
    int grid_nb = 1, grid_cap = 0, grid_it = 2, big_nb = 1, big_cap = 0;
    int *big_cells = nullptr, *d_big_count = nullptr;  // d_big_count[0]: big cells, [1]: huge-cell slices
    int2 *huge_items = nullptr;
    int *move_big_list = nullptr, *d_move_big_count = nullptr;  // cells above move_big_min entries (k_move_big)
    int move_big_min = 4096;                                    // PARSIM_MOVE_BIG_MIN (setting it forces the path)
    bool move_big_force = false;
    int *arr_cnt = nullptr;
    ArrRec *arr_rec = nullptr;
    bool have_newcell = false, have_flags = false;

    int *count = nullptr, *tile_sums = nullptr;
    int *newcell = nullptr, *partner = nullptr;
    int *sort_hist = nullptr, *sort_hist_scan = nullptr;  // radix-sort tile histograms (step 0)
    int *arr_head = nullptr, *arr_next = nullptr, *guest_src = nullptr;
    int *mv_head = nullptr, *dm_head = nullptr, *mv_next = nullptr;  // end-of-step move / dead bags
    unsigned char *fate = nullptr;
    unsigned *tmp_id = nullptr;
    unsigned char *flags = nullptr;
    double2 *com_xy = nullptr;
    double *com_m = nullptr;
    double2 *dbg_acc = nullptr;

    // tombstones (dead CSR) and graveyard
    int *dead_count = nullptr, *dead_new_count = nullptr;
    int *dead_start_cur = nullptr, *dead_start_alt = nullptr;
    unsigned *dead_id_cur = nullptr, *dead_id_alt = nullptr;
    int n_tomb = 0;  // entries in the dead CSR of this strip
    double2 *grave_pos = nullptr, *grave_vel = nullptr;
    unsigned *grave_id = nullptr;
    int *grave_cell = nullptr;
    int n_grave = 0;

    Counters *d_ctr = nullptr, *h_ctr = nullptr;
    // steps enqueued without waiting for the previous one (single strip, parsim_step.inl rank_run)
    bool deferred = false, defer_steps = true;  // defer_steps: PARSIM_SYNC_STEPS=1 switches it off
    Counters *h_ring = nullptr;                 // pinned, kCounterRing entries
    cudaEvent_t ev_ring[kCounterRing] = {nullptr, nullptr};
    long long ring_pos = 0;
    // Small problems are launch-bound (BASELINE configs[1]: about 20 launches per 100 us step): a
    // deferred step is captured once per (buffer parity, kernel choice) as a CUDA graph and replayed.
    struct StepGraph { unsigned long long key; int n_cap; cudaGraphExec_t exec; bool com_after; };
    std::vector<StepGraph> graphs;
    bool use_graphs = false, capturing = false;
    double2 *d_p0 = nullptr;
    int *d_found = nullptr;

    // upload staging
    // two staging sets: the copy engine fills one (copy_stream) while k_ingest reads the other
    double *stage[2][5] = {{nullptr, nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr, nullptr}};
    unsigned *stage_ids[2] = {nullptr, nullptr};
    int stage_cap = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_ingested[2] = {nullptr, nullptr};
    long long stage_pieces = 0;  // staging passes since upload_begin (set = pass & 1)
    long long n_uploaded = 0;
    // the uploaded state (one 64-byte record per slot, parsim_kernels.cuh); the step-0 sort reads
    // it and never modifies it, so parsim_reset only rewinds the bookkeeping
    InitRec *rec0 = nullptr;
    int n0 = 0;                 // records in rec0
    bool slots_id_ordered = true;  // slot order == id order (single strip); else ids are sorted too
    bool opt_keep_initial = false; // accepted for compatibility: reset is always possible now
    bool uploading = false;
    // pieces of the current upload that were sorted as they arrived (single strip, ids in order)
    bool pieces_ok = false, piece_sorted = false, opt_piece_sort = true;
    std::vector<int> piece_first;         // first record of every sorted piece
    int *cs_pieces = nullptr;             // per piece: cell_start[ncells + 1] relative to the piece
    int *d_piece_first = nullptr;
    int cs_pieces_cap = 0;                // pieces cs_pieces has room for
    long long stage_piece_len = 1 << 23;  // particles per staging pass (PARSIM_STAGE_PIECE, tests)

    // multi-strip exchange (parsim_exchange.cuh)
    NcclComm comm;
    LocalBus *bus = nullptr;      // non-null: device-copy transport inside one process, no NCCL
    cudaEvent_t ev_bus = nullptr; // "my send buffers / boundary rows are ready" (local transport)
    int sm_count = 148;           // multiprocessors of this device (persistent grids)
    int up = 0, dn = 0;          // neighbour ranks (periodic)
    int mig_cap = 0;             // records per direction
    double *mig_send[2] = {nullptr, nullptr}, *mig_recv[2] = {nullptr, nullptr};
    int *mig_hdr_send = nullptr, *mig_hdr_recv = nullptr, *h_mig_hdr = nullptr;
    // far records: particles that a close encounter threw beyond the neighbouring row (any strip)
    double *far_send = nullptr, *far_recv = nullptr;
    int *mig_list = nullptr;                // indices of this step's emigrants (k_classify) / late records (k_resolve)
    int mig_list_cap = 0;
    int eager_recs = 8192, far_eager = 32;  // record slots of the first exchange round (<= kEagerRecs / kFarEager)
    double *far_eager_recv = nullptr;       // first-round all-gather: (1 + kFarEager) records per strip
    unsigned long long *h_hdr64 = nullptr;  // pinned: received headers (up, down, far of every strip)

    cudaStream_t stream = nullptr;
    std::string err;
    std::string soft_err;  // a failure only this strip knows about, until the next exchange tells the others

    // options / stats
    bool opt_stats = false, opt_sync_check = false;
    double capacity_factor = 1.35;
    parsim_stats_t stats{};
    std::vector<TimedLaunch> pending;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> event_pool;
    cudaEvent_t run_a = nullptr, run_b = nullptr;
    long long steps_done = 0;
    unsigned long long collisions_total = 0;  // host mirror of d_ctr->collisions (this strip)

    int fail(const std::string &what) {
        err = what;
        return 1;
    }
};

#define PS_CUDA(c, call)                                                                        \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess)                                                                 \
            return (c)->fail(std::string(#call) + ": " + cudaGetErrorString(e__));              \
    } while (0)

inline int grid_for(long long n, int threads = kThreads) {
    long long b = (n + threads - 1) / threads;
    return (int)(b > 0 ? b : 1);
}

// ---- stats: CUDA events around every launch, on the launching stream -----------------------
inline void timed_begin(Rank *c, int k, double bytes) {
    if (!c->opt_stats) return;
    std::pair<cudaEvent_t, cudaEvent_t> ev;
    if (!c->event_pool.empty()) {
        ev = c->event_pool.back();
        c->event_pool.pop_back();
    } else {
        cudaEventCreate(&ev.first);
        cudaEventCreate(&ev.second);
    }
    cudaEventRecord(ev.first, c->stream);
    c->pending.push_back({ev.first, ev.second, k});
    c->stats.kernel_launches[k]++;
    c->stats.kernel_bytes[k] += bytes;
}
inline void timed_end(Rank *c) {
    if (!c->opt_stats) return;
    cudaEventRecord(c->pending.back().b, c->stream);
}
inline int timed_collect(Rank *c) {  // stream must be idle
    for (auto &t : c->pending) {
        float ms = 0;
        PS_CUDA(c, cudaEventElapsedTime(&ms, t.a, t.b));
        c->stats.kernel_ms[t.k] += ms;
        c->event_pool.push_back({t.a, t.b});
    }
    c->pending.clear();
    return 0;
}

inline int check_launch(Rank *c, const char *name) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return c->fail(std::string(name) + " launch: " + cudaGetErrorString(e));
    if (c->opt_sync_check) {
        e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) return c->fail(std::string(name) + ": " + cudaGetErrorString(e));
    }
    return 0;
}

#define PS_LAUNCH(c, kid, bytes, name, ...)        \
    do {                                           \
        timed_begin(c, kid, bytes);                \
        __VA_ARGS__;                               \
        timed_end(c);                              \
        if (check_launch(c, name)) return 1;       \
    } while (0)

// rank r of nranks owns grid rows [r*nc/R, (r+1)*nc/R)
inline void strip_rows(long ncside, int nranks, int rank, long *first_row, long *n_rows) {
    const long r0 = (long)((long long)rank * ncside / nranks);
    const long r1 = (long)((long long)(rank + 1) * ncside / nranks);
    *first_row = r0;
    *n_rows = r1 - r0;
}

// Equal-PARTICLE strips: strip boundaries from the prefix sum of the per-row
// particle counts, so that clustered inputs (negative seeds: normal around the centre) load the
// GPUs evenly.  first_rows[0] = 0 < first_rows[1] < ... < first_rows[nranks] = ncside; every strip
// gets at least one row.  Boundary r is the row where the running count first reaches r/R of
// the total, moved as little as needed to keep the strips non-empty.
inline void balance_rows(long ncside, int nranks, const long long *row_counts, long *first_rows) {
    long long total = 0;
    for (long r = 0; r < ncside; r++) total += row_counts[r];
    first_rows[0] = 0;
    long row = 0;
    long long run = 0;
    for (int k = 1; k < nranks; k++) {
        // smallest row count whose prefix sum is >= k/R of the total
        const long long want = (total * k + nranks - 1) / nranks;
        while (row < ncside && run < want) run += row_counts[row++];
        long b = row;
        const long lo = first_rows[k - 1] + 1, hi = ncside - (nranks - k);
        if (b < lo) b = lo;
        if (b > hi) b = hi;
        while (row < b) run += row_counts[row++];
        first_rows[k] = b;
    }
    first_rows[nranks] = ncside;
}

// first_rows == nullptr: equal-row strips (strip_rows)
int rank_create(Rank *c, double side, long ncside, long long n_part, int rank, int nranks,
                int device, const unsigned char *nccl_id, const long *first_rows = nullptr,
                LocalBus *bus = nullptr);
void rank_destroy(Rank *c);
int rank_upload_begin(Rank *c);
int rank_upload_chunk(Rank *c, long long id_base, const unsigned *ids, long long count, const double *x,
                      const double *y, const double *vx, const double *vy, const double *m);
int rank_stage_alloc(Rank *c);
int rank_ingest_piece(Rank *c, long long id_base, const unsigned *dids, int cnt, double *const st[5], int *up_slot = nullptr);
int rank_reset(Rank *c);
int rank_download(Rank *c, long long capacity, unsigned *ids, double *x, double *y, double *vx,
                  double *vy, double *m, long long *count_out);
int rank_fp64_peak(Rank *c, double *tflops);
int rank_pair_peak(Rank *c, double *pairs_per_s);
int rank_upload_end(Rank *c);
int rank_upload_strip(Rank *c, long long count, const unsigned *ids, const double *x, const double *y,
                      const double *vx, const double *vy, const double *m);
int rank_init_device(Rank *c, long seed);
int rank_stage(Rank *c, int stage);
int rank_ensure_sorted(Rank *c);
int rank_run(Rank *c, long n_steps);
int rank_result_local(Rank *c, double *x0, double *y0, int *found, long long *collisions);
int rank_dump(Rank *c, double *x, double *y, double *vx, double *vy, double *m, int *alive, int *cell,
              int *list_cell, unsigned long long *list_rank, unsigned char *owned);
int rank_dump_cells(Rank *c, double *mx, double *my, double *m);

}  // namespace parsim
