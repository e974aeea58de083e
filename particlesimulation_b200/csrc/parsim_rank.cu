// parsim_rank.cu -- one row strip on one GPU: allocation, upload, the step driver, dumps.
#include "parsim_rank.cuh"

#include <math.h>
#include <stdio.h>
#include <stddef.h>
#include <string.h>

#include <algorithm>

#include "parsim_exchange.cuh"

namespace parsim {

static int device_scan(Rank *c, const int *in, int N, int *out, int *total2);  // parsim_step.inl
static int sort_piece(Rank *c, int first, int cnt, int piece);                  // parsim_step.inl
static int upload_prestep(Rank *c);                                              // parsim_step.inl
static int upload_prestep_strip(Rank *c, bool late, int **where_out);            // parsim_step.inl
int rank_settle_rec_vel(Rank *c);

template <typename T>
static int dev_alloc(Rank *c, T **p, size_t count, bool zero = false) {
    size_t bytes = sizeof(T) * (count ? count : 1);
    cudaError_t e = cudaMalloc((void **)p, bytes);
    if (e != cudaSuccess)
        return c->fail("cudaMalloc(" + std::to_string(bytes) + " B): " + cudaGetErrorString(e));
    if (zero) {
        e = cudaMemset(*p, 0, bytes);
        if (e != cudaSuccess) return c->fail(std::string("cudaMemset: ") + cudaGetErrorString(e));
    }
    return 0;
}
#define PS_ALLOC(c, p, count, zero) \
    do { if (dev_alloc(c, &(p), (size_t)(count), zero)) return 1; } while (0)

// smallest double t with sqrt(t) >= EPSILON: sqrt(s) < EPSILON <=> s < t (sqrt is monotonic and
// correctly rounded), so the collision test needs no square root per pair.
static double collision_threshold2() {
    double t = kEpsilon * kEpsilon;
    while (sqrt(t) >= kEpsilon) t = nextafter(t, 0.0);
    while (sqrt(t) < kEpsilon) t = nextafter(t, INFINITY);
    return t;
}

int rank_create(Rank *c, double side, long ncside, long long n_part, int rank, int nranks,
                int device, const unsigned char *nccl_id, const long *first_rows, LocalBus *bus) {
    if (!(side > 0)) return c->fail("side_length must be > 0");
    if (ncside < 1) return c->fail("grid_size must be >= 1");
    if (n_part < 1) return c->fail("n_particles must be >= 1");
    if ((long long)ncside * ncside > 2000000000LL) return c->fail("grid_size^2 must fit an int (as in the reference)");
    if (n_part > 4000000000LL) return c->fail("n_particles must fit 32-bit ids");
    if (nranks < 1 || rank < 0 || rank >= nranks) return c->fail("bad rank / nranks");
    if (nranks > ncside) return c->fail("more GPUs than grid rows");
    c->side = side; c->nc = ncside; c->n_part = n_part;
    c->rank = rank; c->nranks = nranks; c->device = device;
    int ndev = 0;
    PS_CUDA(c, cudaGetDeviceCount(&ndev));
    if (ndev < 1) return c->fail("no CUDA device: parsim_b200 has no CPU fallback");
    if (device < 0 || device >= ndev) return c->fail("CUDA device index out of range");
    PS_CUDA(c, cudaSetDevice(device));
    PS_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    PS_CUDA(c, cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    c->bus = nranks > 1 ? bus : nullptr;

    Geom &g = c->g;
    g.side = side;
    g.no_strays = 0;
    g.cellw = side / (double)ncside;  // side_length / grid_size with grid_size a long, :160
    g.inv_cellw = 1.0 / g.cellw;
    g.nc = (int)ncside;
    long first_row, n_rows;
    if (first_rows) {
        if (first_rows[0] != 0 || first_rows[nranks] != ncside) return c->fail("strip rows must cover [0, grid_size)");
        for (int r = 0; r < nranks; r++)
            if (first_rows[r + 1] <= first_rows[r]) return c->fail("every strip needs at least one row, in rank order");
        first_row = first_rows[rank];
        n_rows = first_rows[rank + 1] - first_row;
    } else {
        strip_rows(ncside, nranks, rank, &first_row, &n_rows);
    }
    g.row0 = (int)first_row;
    g.rows = (int)n_rows;
    g.ncells = g.rows * g.nc;
    g.cell_base = g.row0 * g.nc;
    c->thr2 = collision_threshold2();
    {   // |float(x) - x| <= side * 2^-24, so an FP32 coordinate difference is off by < side * 2.5e-7
        double delta = side * 2.5e-7;
        double t = (kEpsilon + 1.5 * delta) * (kEpsilon + 1.5 * delta) * (1.0 + 1e-5);
        c->thrf = nextafterf((float)t, INFINITY);
    }
    double ppc = (double)n_part / ((double)ncside * (double)ncside);
    c->com_lanes = ppc <= 6 ? 4 : ppc <= 12 ? 8 : ppc <= 24 ? 16 : 32;

    {   // sub-grid of the collision search: pitch >= EPSILON (with a margin for the binning arithmetic)
        const double hmin = kEpsilon * (1.0 + 1e-6);
        const double fit = floor(g.cellw / hmin);
        const int nb_geom = fit >= 1.0 ? (fit > 1024.0 ? 1024 : (int)fit) : 1;
        c->grid_nb = std::min(nb_geom, 16);
        c->big_nb = std::min(nb_geom, 32);
        // bin-time entries per lane of the warp-per-cell kernel: covers ppc + 6 sigma (Poisson)
        const double want = ppc + 6.0 * sqrt(ppc) + 8.0;
        c->grid_it = want <= 64 ? 2 : want <= 128 ? 4 : want <= 192 ? 6 : want <= 256 ? 8 : 12;
        c->grid_cap = 32 * c->grid_it + 32;
        c->big_cap = 6400;
        c->collide_grid = c->grid_nb >= 3 && ppc >= 16.0;
        if (const char *e = getenv("PARSIM_COLLIDE")) {
            if (!strcmp(e, "pairs")) c->collide_grid = false;
            else if (!strcmp(e, "grid")) c->collide_grid = true;
        }
        if (c->collide_grid) {
#define PS_COLLIDE_ATTR(IT)                                                                              \
    PS_CUDA(c, cudaFuncSetAttribute(k_collide_cells<IT>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                    (int)(kCollideWarps * warp_staging_bytes(32 * IT + 32))))
            PS_COLLIDE_ATTR(1); PS_COLLIDE_ATTR(2); PS_COLLIDE_ATTR(3); PS_COLLIDE_ATTR(4);
            PS_COLLIDE_ATTR(6); PS_COLLIDE_ATTR(8); PS_COLLIDE_ATTR(12);
#undef PS_COLLIDE_ATTR
            PS_CUDA(c, cudaFuncSetAttribute(k_collide_big, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)staging_bytes(c->big_nb * c->big_nb, c->big_cap)));
        }
    }
    if (const char *e = getenv("PARSIM_FORCE_VARIANT")) c->force_variant = atoi(e);
    if (const char *e = getenv("PARSIM_SPARSE_VARIANT")) c->sparse_variant = atoi(e) == 1 ? 1 : 5;
    {   // only instantiations that the variant-agreement test covers (tests/test_gpu_parity.py)
        const int v = c->force_variant;
        if (!(v >= 0 && v <= 6) && v != 10 && v != 11)
            return c->fail("PARSIM_FORCE_VARIANT: unknown force-kernel variant " + std::to_string(v));
    }
    if (const char *e = getenv("PARSIM_CAPACITY_PCT")) {
        double pct = atof(e);
        if (pct >= 100) c->capacity_factor = pct / 100.0;
    }
    long long cap;
    if (nranks == 1) {
        cap = n_part;
        c->mig_cap = 0;
    } else {
        long long per_row = n_part / ncside + 1;
        long long mc = std::max<long long>(65536, 2 * per_row);
        c->mig_cap = (int)std::min<long long>(mc, n_part);
        if (const char *e = getenv("PARSIM_MIG_CAP")) c->mig_cap = std::max(1, atoi(e));  // tests: force an overflow
        cap = (long long)((double)n_part / nranks * c->capacity_factor) + 4LL * c->mig_cap + 1024 +
              2LL * nranks * kFarCap;
        cap = std::min<long long>(cap, n_part + 2LL * c->mig_cap + 2LL * nranks * kFarCap);
    }
    if (cap > 2147480000LL) return c->fail("strip too large for 32-bit indices: use more GPUs");
    c->cap = (int)cap;

    const size_t N = (size_t)c->cap, C = (size_t)g.ncells;
    PS_ALLOC(c, c->rec0, N, false);
    PS_ALLOC(c, c->pos_cur, N, false); PS_ALLOC(c, c->pos_alt, N, false);
    PS_ALLOC(c, c->vel_cur, N, false); PS_ALLOC(c, c->vel_alt, N, false);
    if (const char *e = getenv("PARSIM_MOVE_BIG_MIN")) {
        c->move_big_min = std::max(1, atoi(e));
        c->move_big_force = true;
    }
    PS_ALLOC(c, c->move_big_list, std::min(C, (N + (size_t)n_part) / (size_t)c->move_big_min + 16), false);
    PS_ALLOC(c, c->d_move_big_count, 1, true);
    if (!c->collide_grid) PS_ALLOC(c, c->posf, N, false);
    else {
        PS_ALLOC(c, c->big_cells, C, false); PS_ALLOC(c, c->d_big_count, 2, true);
        PS_ALLOC(c, c->huge_items, N / kThreads + 2 * C + 16, false);
        PS_ALLOC(c, c->arr_cnt, C, true); PS_ALLOC(c, c->arr_rec, C * kArrSlots, false);
    }
    PS_ALLOC(c, c->m_cur, N, false);   PS_ALLOC(c, c->m_alt, N, false);
    PS_ALLOC(c, c->id_cur, N, false);  PS_ALLOC(c, c->id_alt, N, false);
    PS_ALLOC(c, c->cell_cur, N, false); PS_ALLOC(c, c->cell_alt, N, false);
    PS_ALLOC(c, c->cs_cur, C + 1, true); PS_ALLOC(c, c->cs_alt, C + 1, true);
    PS_ALLOC(c, c->count, C, true);
    const size_t H = (size_t)kSortMaxBins * (N / kSortTile + 1);  // radix-sort tile histograms
    PS_ALLOC(c, c->tile_sums, 2 * (std::max(C, H) / kScanTile + 2), true);
    PS_ALLOC(c, c->sort_hist, H, false); PS_ALLOC(c, c->sort_hist_scan, H + 1, false);
    PS_ALLOC(c, c->newcell, N, false);
    PS_ALLOC(c, c->partner, N, false);
    PS_ALLOC(c, c->tmp_id, N, false);
    PS_ALLOC(c, c->arr_head, C, false); PS_ALLOC(c, c->arr_next, N, false);
    PS_ALLOC(c, c->mv_head, C, false); PS_ALLOC(c, c->dm_head, C, false); PS_ALLOC(c, c->mv_next, N, false);
    PS_ALLOC(c, c->fate, N, true);
    PS_ALLOC(c, c->flags, N, true);
    PS_ALLOC(c, c->com_xy, (size_t)(g.rows + 2) * g.nc, true);
    PS_ALLOC(c, c->com_m, (size_t)(g.rows + 2) * g.nc, true);
    PS_ALLOC(c, c->dead_count, C, true); PS_ALLOC(c, c->dead_new_count, C, true);
    PS_ALLOC(c, c->dead_start_cur, C + 1, true); PS_ALLOC(c, c->dead_start_alt, C + 1, true);
    PS_ALLOC(c, c->dead_id_cur, N, false); PS_ALLOC(c, c->dead_id_alt, N, false);
    PS_ALLOC(c, c->grave_pos, N, false); PS_ALLOC(c, c->grave_vel, N, false);
    PS_ALLOC(c, c->grave_id, N, false); PS_ALLOC(c, c->grave_cell, N, false);
    PS_ALLOC(c, c->d_ctr, 1, true);
    PS_ALLOC(c, c->d_p0, 1, true);
    PS_ALLOC(c, c->d_found, 1, true);
    PS_CUDA(c, cudaMallocHost((void **)&c->h_ctr, sizeof(Counters)));
    memset(c->h_ctr, 0, sizeof(Counters));
    PS_CUDA(c, cudaMallocHost((void **)&c->h_ring, kCounterRing * sizeof(Counters)));
    for (int k = 0; k < kCounterRing; k++) PS_CUDA(c, cudaEventCreateWithFlags(&c->ev_ring[k], cudaEventDisableTiming));
    c->defer_steps = c->use_graphs = n_part <= 4000000;  // above that a step is milliseconds of kernels: nothing to gain
    if (const char *e = getenv("PARSIM_SYNC_STEPS")) c->defer_steps = atoi(e) == 0;
    if (const char *e = getenv("PARSIM_GRAPHS")) c->use_graphs = atoi(e) != 0;
    if (const char *e = getenv("PARSIM_PIECE_SORT")) c->opt_piece_sort = atoi(e) != 0;
    if (const char *e = getenv("PARSIM_FUSE_COM")) c->fuse_com = atoi(e) != 0;
    if (const char *e = getenv("PARSIM_VEL_LATE")) c->opt_vel_late = atoi(e) != 0;
    if (const char *e = getenv("PARSIM_STAGE_PIECE")) c->stage_piece_len = std::max(64, atoi(e));
    PS_CUDA(c, cudaEventCreate(&c->run_a));
    PS_CUDA(c, cudaEventCreate(&c->run_b));
    if (nranks > 1) {
        if (exchange_create(c, nccl_id)) return 1;
    }
    PS_CUDA(c, cudaDeviceSynchronize());  // the zero-fills above ran on the legacy stream
    return 0;
}

void rank_destroy(Rank *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    exchange_destroy(c);
    void *ptrs[] = {c->pos_cur, c->pos_alt, c->vel_cur, c->vel_alt, c->m_cur, c->m_alt, c->id_cur,
                    c->id_alt, c->cell_cur, c->cell_alt, c->cs_cur, c->cs_alt, c->count, c->tile_sums,
                    c->sort_hist, c->sort_hist_scan, c->newcell, c->partner, c->tmp_id, c->arr_head, c->arr_next,
                    c->flags, c->com_xy, c->com_m, c->dbg_acc, c->dead_count, c->dead_new_count,
                    c->dead_start_cur, c->dead_start_alt, c->dead_id_cur, c->dead_id_alt,
                    c->mv_head, c->dm_head, c->mv_next, c->fate, c->grave_pos, c->grave_vel, c->grave_id,
                    c->grave_cell, c->d_ctr, c->d_p0, c->d_found, c->guest_src, c->posf, c->rec0,
                    c->big_cells, c->d_big_count, c->huge_items, c->move_big_list, c->d_move_big_count, c->arr_cnt, c->arr_rec,
                    c->cs_pieces, c->d_piece_first,
                    c->stage_ids[0], c->stage_ids[1], c->stage[0][0], c->stage[0][1], c->stage[0][2],
                    c->stage[0][3], c->stage[0][4], c->stage[1][0], c->stage[1][1], c->stage[1][2],
                    c->stage[1][3], c->stage[1][4]};
    for (void *p : ptrs) if (p) cudaFree(p);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    for (int b = 0; b < 2; b++) {
        if (c->ev_copied[b]) cudaEventDestroy(c->ev_copied[b]);
        if (c->ev_ingested[b]) cudaEventDestroy(c->ev_ingested[b]);
    }
    if (c->h_ctr) cudaFreeHost(c->h_ctr);
    for (auto &sg : c->graphs) cudaGraphExecDestroy(sg.exec);
    c->graphs.clear();
    if (c->up_slot) cudaFree(c->up_slot);
    if (c->late_vx) cudaFree(c->late_vx);
    if (c->late_vy) cudaFree(c->late_vy);
    if (c->h_ring) cudaFreeHost(c->h_ring);
    for (int k = 0; k < kCounterRing; k++) if (c->ev_ring[k]) cudaEventDestroy(c->ev_ring[k]);
    for (auto &t : c->pending) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    for (auto &e : c->event_pool) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    if (c->run_a) cudaEventDestroy(c->run_a);
    if (c->run_b) cudaEventDestroy(c->run_b);
    if (c->stream) cudaStreamDestroy(c->stream);
}

// ---- upload -------------------------------------------------------------------------------

static int reset_state(Rank *c) {
    c->n_uploaded = 0;
    const size_t C = (size_t)c->g.ncells;
    c->n_alive = 0; c->n_guest = 0; c->sorted = false; c->com_valid = false; c->count_pending = false; c->acc0 = nullptr; c->next_stage = 0;
    c->rec_vel_pending = false; c->halo_valid = false; c->late_count = 0;
    c->piece_sorted = false;
    c->n_tomb = 0; c->n_grave = 0; c->steps_done = 0; c->collisions_total = 0;
    c->have_newcell = c->have_flags = false;
    c->pos_latest = c->pos_cur;
    c->soft_err.clear();
    // everything but the four statistics counters (parsim_stats_reset clears those)
    PS_CUDA(c, cudaMemsetAsync(&c->d_ctr->collisions, 0, sizeof(unsigned long long), c->stream));
    PS_CUDA(c, cudaMemsetAsync(&c->d_ctr->deaths_step, 0, sizeof(Counters) - offsetof(Counters, deaths_step),
                               c->stream));
    PS_CUDA(c, cudaMemsetAsync(c->count, 0, sizeof(int) * C, c->stream));
    PS_CUDA(c, cudaMemsetAsync(c->dead_count, 0, sizeof(int) * C, c->stream));
    PS_CUDA(c, cudaMemsetAsync(c->dead_new_count, 0, sizeof(int) * C, c->stream));
    PS_CUDA(c, cudaMemsetAsync(c->dead_start_cur, 0, sizeof(int) * (C + 1), c->stream));
    PS_CUDA(c, cudaMemsetAsync(c->arr_head, 0xFF, sizeof(int) * C, c->stream));  // empty bags (-1)
    if (c->arr_cnt) PS_CUDA(c, cudaMemsetAsync(c->arr_cnt, 0, sizeof(int) * C, c->stream));
    PS_CUDA(c, cudaMemsetAsync(c->mv_head, 0xFF, sizeof(int) * C, c->stream));
    PS_CUDA(c, cudaMemsetAsync(c->dm_head, 0xFF, sizeof(int) * C, c->stream));
    return 0;
}

// An upload with late velocities leaves the records' velocity fields for later (late_vx / late_vy
// hold them in slot order): whoever reads the records' velocities calls this first.
int rank_settle_rec_vel(Rank *c) {
    if (!c->rec_vel_pending) return 0;
    c->rec_vel_pending = false;
    if (c->late_count > 0) {  // strip upload: upload order, through up_slot
        const int n = (int)c->late_count;
        k_rec_vel_strip<<<grid_for(n), kThreads, 0, c->stream>>>(n, c->up_slot, c->late_vx, c->late_vy, c->rec0);
    } else {
        const int n = (int)c->n_part;
        k_rec_vel<<<grid_for(n), kThreads, 0, c->stream>>>(n, c->late_vx, c->late_vy, c->rec0);
    }
    return check_launch(c, "k_rec_vel");
}

static int late_alloc(Rank *c, size_t count) {
    if (count <= c->late_cap) return 0;
    if (c->late_vx) cudaFree(c->late_vx);
    if (c->late_vy) cudaFree(c->late_vy);
    c->late_vx = c->late_vy = nullptr;
    c->late_cap = 0;
    PS_ALLOC(c, c->late_vx, count, false);
    PS_ALLOC(c, c->late_vy, count, false);
    c->late_cap = count;
    return 0;
}

int rank_upload_begin(Rank *c) {
    PS_CUDA(c, cudaSetDevice(c->device));
    PS_CUDA(c, cudaStreamSynchronize(c->stream));  // an abandoned upload may still be reading the staging sets
    if (reset_state(c)) return 1;
    c->uploading = true;
    c->stage_pieces = 0;
    c->piece_first.clear();
    c->pieces_ok = false;
    if (c->nranks == 1 && c->opt_piece_sort) {
        // room for the pieces' cell_start tables (at most 64 pieces and 1 GiB: else the full sort at step 0)
        const long long pieces = std::min<long long>(64, (c->n_part + c->stage_piece_len - 1) / c->stage_piece_len + 1);
        const size_t need = (size_t)pieces * ((size_t)c->g.ncells + 1);
        if (!c->cs_pieces && need * sizeof(int) <= ((size_t)1 << 30)) {
            PS_ALLOC(c, c->cs_pieces, need, false);
            PS_ALLOC(c, c->d_piece_first, 64, false);
            c->cs_pieces_cap = (int)pieces;
        }
        c->pieces_ok = c->cs_pieces != nullptr;
    }
    return 0;
}

// Staging passes of 8 Mi particles (5 x 64 MiB per set).  The host->device copies of a pass run
// on copy_stream into one staging set while the ingest kernels (on the step stream) interleave the
// previous pass out of the other set, so the copy engine never waits for a kernel.
int rank_stage_alloc(Rank *c) {
    if (c->stage[0][0]) return 0;
    PS_CUDA(c, cudaSetDevice(c->device));
    c->stage_cap = (int)std::min<long long>(c->stage_piece_len, c->n_part);
    for (int b = 0; b < 2; b++) {
        for (int k = 0; k < 5; k++) PS_ALLOC(c, c->stage[b][k], c->stage_cap, false);
        PS_ALLOC(c, c->stage_ids[b], c->stage_cap, false);
        PS_CUDA(c, cudaEventCreateWithFlags(&c->ev_copied[b], cudaEventDisableTiming));
        PS_CUDA(c, cudaEventCreateWithFlags(&c->ev_ingested[b], cudaEventDisableTiming));
    }
    PS_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    return 0;
}

// `cnt` staged particles (device SoA `st`, ids `dids` or id_base + index) into this strip's records.
// The staging arrays may live on another GPU of the process (peer access, group upload).
int rank_ingest_piece(Rank *c, long long id_base, const unsigned *dids, int cnt, double *const st[5], int *up_slot) {
    if (c->nranks == 1) {
        PS_LAUNCH(c, K_INGEST, 84.0 * cnt, "k_ingest",
                  k_ingest<<<grid_for(cnt), kThreads, 0, c->stream>>>(
                      c->g, id_base, dids, cnt, st[0], st[1], st[2], st[3], st[4], c->cap, c->rec0, c->d_ctr));
    } else {  // stable compaction of this strip's particles: count, scan, write
        const int nb = grid_for(cnt);
        PS_LAUNCH(c, K_INGEST, 8.0 * cnt, "k_ingest_strip<0>",
                  k_ingest_strip<0><<<nb, kThreads, 0, c->stream>>>(
                      c->g, id_base, dids, cnt, st[0], st[1], st[2], st[3], st[4],
                      c->cap - 2 * c->mig_cap, c->sort_hist, c->rec0, c->d_ctr));
        if (device_scan(c, c->sort_hist, nb, c->sort_hist_scan, nullptr)) return 1;
        PS_LAUNCH(c, K_INGEST, 84.0 * cnt, "k_ingest_strip<1>",
                  k_ingest_strip<1><<<nb, kThreads, 0, c->stream>>>(
                      c->g, id_base, dids, cnt, st[0], st[1], st[2], st[3], st[4],
                      c->cap - 2 * c->mig_cap, c->sort_hist_scan, c->rec0, c->d_ctr, up_slot));
        k_advance_staged<<<1, 1, 0, c->stream>>>(c->d_ctr, c->sort_hist_scan + nb);
    }
    return 0;
}

int rank_upload_chunk(Rank *c, long long id_base, const unsigned *ids, long long count, const double *x,
                      const double *y, const double *vx, const double *vy, const double *m) {
    if (!c->uploading) return c->fail("parsim_upload_chunk outside upload_begin/upload_end");
    if (count < 0 || (!ids && (id_base < 0 || id_base + count > c->n_part)))
        return c->fail("upload chunk outside [0, n_part)");
    PS_CUDA(c, cudaSetDevice(c->device));
    if (rank_stage_alloc(c)) return 1;
    const double *src[5] = {x, y, vx, vy, m};
    // Late velocities: a single strip that gets ALL its particles in this one call (parsim_upload) and
    // sorts the pieces as they arrive copies x, y, m of every piece first and the two velocity arrays
    // last, and runs the part of the first timestep that needs no velocity -- the merge of the pieces,
    // the centres of mass, the force sums -- while those 16 bytes per particle are still on the link.
    const bool vel_late = c->opt_vel_late && c->nranks == 1 && !ids && id_base == 0 && count == c->n_part &&
                          c->n_uploaded == 0 && c->pieces_ok && count > c->stage_cap && c->force_variant <= 0 &&
                          !c->opt_stats && !c->opt_sync_check && !c->dbg_acc;
    for (long long off = 0; off < count; off += c->stage_cap) {
        const int cnt = (int)std::min<long long>(c->stage_cap, count - off);
        const int b = (int)(c->stage_pieces & 1);
        double *const *st = c->stage[b];
        // the set is free once the k_ingest of two passes ago has read it
        if (c->stage_pieces >= 2) PS_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_ingested[b], 0));
        for (int k = 0; k < 5; k++)
            if (!vel_late || (k != 2 && k != 3))
                PS_CUDA(c, cudaMemcpyAsync(st[k], src[k] + off, sizeof(double) * cnt, cudaMemcpyHostToDevice,
                                           c->copy_stream));
        if (ids)
            PS_CUDA(c, cudaMemcpyAsync(c->stage_ids[b], ids + off, sizeof(unsigned) * cnt,
                                       cudaMemcpyHostToDevice, c->copy_stream));
        PS_CUDA(c, cudaEventRecord(c->ev_copied[b], c->copy_stream));
        PS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[b], 0));
        double *const late_set[5] = {st[0], st[1], nullptr, nullptr, st[4]};
        if (rank_ingest_piece(c, id_base + off, ids ? c->stage_ids[b] : nullptr, cnt, vel_late ? late_set : st)) return 1;
        PS_CUDA(c, cudaEventRecord(c->ev_ingested[b], c->stream));
        c->stage_pieces++;
        // sort the piece now, under the shadow of the next piece's copy (parsim_step.inl sort_piece):
        // only while the pieces arrive in id order, one after the other, and their tables fit
        if (c->pieces_ok && (ids || id_base + off != c->n_uploaded + off)) c->pieces_ok = false;
        if (c->pieces_ok && (int)c->piece_first.size() >= c->cs_pieces_cap) c->pieces_ok = false;
        if (c->pieces_ok) {
            const int piece = (int)c->piece_first.size();
            c->piece_first.push_back((int)(id_base + off));
            if (sort_piece(c, (int)(id_base + off), cnt, piece)) return 1;
        }
    }
    if (vel_late) {
        if (late_alloc(c, (size_t)c->n_part)) return 1;
        c->n_uploaded = count;
        const bool pre = c->pieces_ok;  // (false: more pieces than tables -- the records are completed, step 0 sorts them)
        int *where = c->mv_next;        // slot -> ordered position; free until the collision phase
        if (pre) {
            c->n_alive = c->n0 = (int)c->n_part;
            c->piece_sorted = true;
            c->slots_id_ordered = true;
            k_set_int<<<1, 1, 0, c->stream>>>(&c->d_ctr->n_next, c->n_alive);
            if (upload_prestep(c)) return 1;
            k_slot_positions<<<grid_for(c->n_alive), kThreads, 0, c->stream>>>(c->n_alive, c->id_cur, where);
        }
        // the velocity arrays, piece by piece: every piece is scattered into the ordered array as soon as
        // it is there, so only the last one is not hidden behind the copy
        for (long long off = 0; off < count; off += c->stage_cap) {
            const int cnt = (int)std::min<long long>(c->stage_cap, count - off);
            PS_CUDA(c, cudaMemcpyAsync(c->late_vx + off, vx + off, sizeof(double) * cnt, cudaMemcpyHostToDevice, c->copy_stream));
            PS_CUDA(c, cudaMemcpyAsync(c->late_vy + off, vy + off, sizeof(double) * cnt, cudaMemcpyHostToDevice, c->copy_stream));
            PS_CUDA(c, cudaEventRecord(c->ev_copied[0], c->copy_stream));
            PS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[0], 0));
            if (pre)
                k_vel_scatter<<<grid_for(cnt), kThreads, 0, c->stream>>>((int)off, cnt, where, c->late_vx, c->late_vy, c->vel_cur);
        }
        c->rec_vel_pending = true;  // the records get theirs when somebody needs them (reset, full sort, download)
        if (!pre && rank_settle_rec_vel(c)) return 1;
        PS_CUDA(c, cudaStreamSynchronize(c->copy_stream));  // the caller's arrays are free again
        return check_launch(c, "k_vel_scatter");
    }
    // the caller's arrays are free again when this returns ("they are copied"); the ingest kernels
    // may still be running -- parsim_upload_end waits for them
    PS_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    c->n_uploaded += count;
    return 0;
}

int rank_upload_end(Rank *c) {
    if (!c->uploading) return c->fail("parsim_upload_end without parsim_upload_begin");
    PS_CUDA(c, cudaSetDevice(c->device));
    c->uploading = false;
    PS_CUDA(c, cudaMemcpyAsync(c->h_ctr, c->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->h_ctr->error == DE_CAPACITY)
        return c->fail("strip capacity exceeded during upload (skewed distribution?): set the environment "
                       "variable PARSIM_CAPACITY_PCT (default 135 = 1.35 x n_part / n_gpus) before creating");
    if (c->nranks > 1) c->n_alive = c->h_ctr->n_staged;
    else if (c->n_uploaded != c->n_part) return c->fail("upload did not cover all n_part particles");
    else c->n_alive = (int)c->n_part;
    c->pos_latest = c->pos_cur;
    c->n0 = c->n_alive;
    // (an upload with late velocities has merged its pieces already: c->sorted)
    c->piece_sorted = !c->sorted && c->pieces_ok && !c->piece_first.empty() && c->n_uploaded == c->n_part;
    k_set_int<<<1, 1, 0, c->stream>>>(&c->d_ctr->n_next, c->n_alive);  // the kernels' particle count
    c->slots_id_ordered = true;
    if (c->nranks > 1 && c->n_alive > 1) {  // explicit-id uploads may come in any order
        PS_CUDA(c, cudaMemsetAsync(c->d_found, 0, sizeof(int), c->stream));
        k_rec_ids_ascending<<<grid_for(c->n_alive), kThreads, 0, c->stream>>>(c->n_alive, c->rec0, c->d_found);
        int unordered = 0;
        PS_CUDA(c, cudaMemcpyAsync(&unordered, c->d_found, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        c->slots_id_ordered = !unordered;
    }
    return 0;
}

// parsim_upload_strip: the whole strip of this rank in ONE call.  On a multi-rank context the call is
// COLLECTIVE (every rank makes it): the strips agree that their uploads went well, bin, take their
// centres of mass and exchange the ghost rows before it returns, so the first timestep starts with
// the force stage.  Where nothing forbids it (instrumentation, explicit kernel variants) x, y, m and
// the ids of every piece travel first and the two velocity arrays last, and the binning, the centres
// of mass, the ghost rows and the force sums of the first timestep run under their copy.
int rank_upload_strip(Rank *c, long long count, const unsigned *ids, const double *x, const double *y,
                      const double *vx, const double *vy, const double *m) {
    if (c->nranks == 1 || c->bus) {  // nothing to agree on: the plain sequence
        if (count < 0 || (count > 0 && (!ids || !x || !y || !vx || !vy || !m))) return c->fail("parsim_upload_strip: NULL array");
        if (rank_upload_begin(c)) return 1;
        if (rank_upload_chunk(c, 0, ids, count, x, y, vx, vy, m)) return 1;
        return rank_upload_end(c);
    }
    // Whatever goes wrong on THIS strip before the agreement (bad arguments, allocation, capacity) must not
    // keep it from the agreement: the other strips would wait for it there.
    bool late = false;
    auto phase_a = [&]() -> int {
    if (count < 0 || (count > 0 && (!ids || !x || !y || !vx || !vy || !m)))
        return c->fail("parsim_upload_strip: NULL array (ids, x, y, vx, vy, m of the strip's particles)");
    if (rank_upload_begin(c)) return 1;
    if (rank_stage_alloc(c)) return 1;
    late = c->opt_vel_late && c->force_variant <= 0 && !c->opt_stats && !c->opt_sync_check && !c->dbg_acc && count > 0;
    if (late) {
        if (late_alloc(c, (size_t)count)) return 1;
        if ((size_t)count > c->up_slot_cap) {
            if (c->up_slot) cudaFree(c->up_slot);
            c->up_slot = nullptr;
            c->up_slot_cap = 0;
            PS_ALLOC(c, c->up_slot, (size_t)count, false);
            c->up_slot_cap = (size_t)count;
        }
    }
    const double *src[5] = {x, y, vx, vy, m};
    for (long long off = 0; off < count; off += c->stage_cap) {
        const int cnt = (int)std::min<long long>(c->stage_cap, count - off);
        const int b = (int)(c->stage_pieces & 1);
        double *const *st = c->stage[b];
        if (c->stage_pieces >= 2) PS_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_ingested[b], 0));
        for (int k = 0; k < 5; k++)
            if (!late || (k != 2 && k != 3))
                PS_CUDA(c, cudaMemcpyAsync(st[k], src[k] + off, sizeof(double) * cnt, cudaMemcpyHostToDevice, c->copy_stream));
        PS_CUDA(c, cudaMemcpyAsync(c->stage_ids[b], ids + off, sizeof(unsigned) * cnt, cudaMemcpyHostToDevice, c->copy_stream));
        PS_CUDA(c, cudaEventRecord(c->ev_copied[b], c->copy_stream));
        PS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[b], 0));
        double *const late_set[5] = {st[0], st[1], nullptr, nullptr, st[4]};
        if (rank_ingest_piece(c, off, c->stage_ids[b], cnt, late ? late_set : st, late ? c->up_slot + off : nullptr)) return 1;
        PS_CUDA(c, cudaEventRecord(c->ev_ingested[b], c->stream));
        c->stage_pieces++;
    }
    c->n_uploaded = count;
    c->pieces_ok = false;
    return rank_upload_end(c);  // the strip's particle count, errors, slot order
    };
    if (phase_a()) exchange_soft_fail(c, c->err);
    c->uploading = false;
    int *where = nullptr;
    if (upload_prestep_strip(c, late, &where)) return 1;
    if (late) {
        for (long long off = 0; off < count; off += c->stage_cap) {
            const int cnt = (int)std::min<long long>(c->stage_cap, count - off);
            PS_CUDA(c, cudaMemcpyAsync(c->late_vx + off, vx + off, sizeof(double) * cnt, cudaMemcpyHostToDevice, c->copy_stream));
            PS_CUDA(c, cudaMemcpyAsync(c->late_vy + off, vy + off, sizeof(double) * cnt, cudaMemcpyHostToDevice, c->copy_stream));
            PS_CUDA(c, cudaEventRecord(c->ev_copied[0], c->copy_stream));
            PS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[0], 0));
            k_vel_scatter_strip<<<grid_for(cnt), kThreads, 0, c->stream>>>((int)off, cnt, c->up_slot, where, c->late_vx,
                                                                           c->late_vy, c->vel_cur);
        }
        c->late_count = count;
        c->rec_vel_pending = true;
        PS_CUDA(c, cudaStreamSynchronize(c->copy_stream));  // the caller's arrays are free again
    }
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    return check_launch(c, "parsim_upload_strip");
}

// Restart from the uploaded state: the records in rec0 are untouched by the step loop.
int rank_reset(Rank *c) {
    PS_CUDA(c, cudaSetDevice(c->device));
    if (c->n0 <= 0 || c->uploading) return c->fail("parsim_reset: nothing uploaded yet");
    const int n0 = c->n0;
    if (rank_settle_rec_vel(c)) return 1;  // (reset_state drops the flag)
    if (reset_state(c)) return 1;
    c->n_alive = c->n0 = n0;
    k_set_int<<<1, 1, 0, c->stream>>>(&c->d_ctr->n_next, c->n_alive);
    c->n_uploaded = c->n_part;
    c->pos_latest = c->pos_cur;
    return 0;
}

// strip particles (alive residents, current order) to host SoA + ids; *count_out = how many
int rank_download(Rank *c, long long capacity, unsigned *ids, double *x, double *y, double *vx,
                  double *vy, double *m, long long *count_out) {
    PS_CUDA(c, cudaSetDevice(c->device));
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    const size_t n = (size_t)c->n_alive;
    *count_out = (long long)n;
    if ((long long)n > capacity) return c->fail("parsim_download: host arrays too small");
    if (!c->sorted) {  // still the uploaded records, slot order
        if (rank_settle_rec_vel(c)) return 1;
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        const size_t piece = 1 << 22;
        std::vector<InitRec> tmp(std::min(n, piece));
        for (size_t off = 0; off < n; off += piece) {
            const size_t cnt = std::min(piece, n - off);
            PS_CUDA(c, cudaMemcpy(tmp.data(), c->rec0 + off, sizeof(InitRec) * cnt, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < cnt; i++) {
                const InitRec &r = tmp[i];
                x[off + i] = r.pos.x; y[off + i] = r.pos.y; vx[off + i] = r.vel.x; vy[off + i] = r.vel.y;
                m[off + i] = r.m; ids[off + i] = r.id;
            }
        }
        return 0;
    }
    std::vector<double2> tmp(n);
    PS_CUDA(c, cudaMemcpy(tmp.data(), c->pos_latest, sizeof(double2) * n, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) { x[i] = tmp[i].x; y[i] = tmp[i].y; }
    PS_CUDA(c, cudaMemcpy(tmp.data(), c->vel_cur, sizeof(double2) * n, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) { vx[i] = tmp[i].x; vy[i] = tmp[i].y; }
    PS_CUDA(c, cudaMemcpy(m, c->m_cur, sizeof(double) * n, cudaMemcpyDeviceToHost));
    PS_CUDA(c, cudaMemcpy(ids, c->id_cur, sizeof(unsigned) * n, cudaMemcpyDeviceToHost));
    return 0;
}

// ---- device-side init_particles --------------------------------------------------------------
static unsigned xs_step_host(unsigned s) { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; }
static unsigned xs_apply_host(const XsMatrix &A, unsigned v) {
    unsigned r = 0;
    for (int b = 0; b < 32; b++) if ((v >> b) & 1u) r ^= A.col[b];
    return r;
}
static XsMatrix xs_mul_host(const XsMatrix &A, const XsMatrix &B) {  // A * B
    XsMatrix R;
    for (int b = 0; b < 32; b++) R.col[b] = xs_apply_host(A, B.col[b]);
    return R;
}

int rank_init_device(Rank *c, long seed) {
    PS_CUDA(c, cudaSetDevice(c->device));
    const int iseed = (int)seed;  // main stores the seed in an int, serial/parsim.cpp:465
    if (iseed < 0)
        return c->fail("parsim_init_device: negative seeds use rejection sampling (serial/parsim.cpp:34-43);"
                       " generate on the host and upload");
    if (reset_state(c)) return 1;
    const bool strips = c->nranks > 1;
    const int blocks = c->sm_count * 8;
    // ids a thread generates are T apart: the whole grid (one strip) or one block (several strips)
    const unsigned long long T = strips ? (unsigned long long)kThreads : (unsigned long long)blocks * kThreads;
    XsJumpTable jt;
    for (int b = 0; b < 32; b++) jt.pow2[0].col[b] = xs_step_host(1u << b);
    for (int j = 1; j < 34; j++) jt.pow2[j] = xs_mul_host(jt.pow2[j - 1], jt.pow2[j - 1]);
    XsMatrix acc;
    for (int b = 0; b < 32; b++) acc.col[b] = 1u << b;  // identity
    unsigned long long e = 5ull * T - 5ull;
    for (int j = 0; e; j++, e >>= 1) if (e & 1ull) acc = xs_mul_host(jt.pow2[j], acc);
    jt.stride = acc;
    XsJumpTable *d_jt = nullptr;
    PS_ALLOC(c, d_jt, 1, false);
    PS_CUDA(c, cudaMemcpyAsync(d_jt, &jt, sizeof(jt), cudaMemcpyHostToDevice, c->stream));
    const unsigned s0 = (unsigned)abs(iseed) + 987654321u;  // :24
    if (!strips) {
        PS_LAUNCH(c, K_INIT, 44.0 * c->n_part, "k_init_particles",
                  k_init_particles<<<blocks, kThreads, 0, c->stream>>>(
                      c->g, d_jt, s0, c->n_part, c->rec0));
    } else {  // every strip walks the whole stream and keeps its rows, in id order
        const int nb = grid_for(c->n_part, kThreads * kInitPer);
        PS_LAUNCH(c, K_INIT, 0.0, "k_init_strip<0>",
                  k_init_strip<0><<<nb, kThreads, 0, c->stream>>>(c->g, d_jt, s0, c->n_part, c->cap - 2 * c->mig_cap,
                                                                 c->sort_hist, c->rec0, c->d_ctr));
        if (device_scan(c, c->sort_hist, nb, c->sort_hist_scan, nullptr)) return 1;
        PS_LAUNCH(c, K_INIT, 64.0 * c->n_part / c->nranks, "k_init_strip<1>",
                  k_init_strip<1><<<nb, kThreads, 0, c->stream>>>(c->g, d_jt, s0, c->n_part, c->cap - 2 * c->mig_cap,
                                                                 c->sort_hist_scan, c->rec0, c->d_ctr));
        k_advance_staged<<<1, 1, 0, c->stream>>>(c->d_ctr, c->sort_hist_scan + nb);
    }
    c->uploading = true;
    c->n_uploaded = c->n_part;
    int rc = rank_upload_end(c);
    cudaFree(d_jt);
    return rc;
}

int rank_fp64_peak(Rank *c, double *tflops) {
    PS_CUDA(c, cudaSetDevice(c->device));
    const int iters = 1 << 15, blocks = c->sm_count * 8;
    double *d = nullptr;
    PS_ALLOC(c, d, 1, true);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(a, c->stream);
        k_fp64_peak<<<blocks, kThreads, 0, c->stream>>>(iters, d);
        cudaEventRecord(b, c->stream);
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(d);
    *tflops = 2.0 * 8.0 * iters * (double)blocks * kThreads / (best * 1e-3) / 1e12;
    return 0;
}

// pair interactions per second of k_pair_peak (the force loop's own instruction mix, no memory)
int rank_pair_peak(Rank *c, double *pairs_per_s) {
    PS_CUDA(c, cudaSetDevice(c->device));
    const int iters = 1 << 13, blocks = c->sm_count * 8;
    double *d = nullptr;
    PS_ALLOC(c, d, 1, true);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(a, c->stream);
        k_pair_peak<<<blocks, kThreads, 0, c->stream>>>(iters, d);
        cudaEventRecord(b, c->stream);
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        if (rep > 0 && ms < best) best = ms;
    }
    *pairs_per_s = 4.0 * iters * (double)blocks * kThreads / (best * 1e-3);
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(d);
    return 0;
}

}  // namespace parsim

#include "parsim_step.inl"
