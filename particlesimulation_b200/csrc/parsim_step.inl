// parsim_step.inl -- the per-step driver (included by parsim_rank.cu).
// Stage numbering follows include/parsim_b200.h (parsim_debug_stage).
namespace parsim {

static bool scan_few_ok() {  // PARSIM_SCAN_FEW=0: always the three-launch scans (A/B)
    static const bool v = !getenv("PARSIM_SCAN_FEW") || atoi(getenv("PARSIM_SCAN_FEW")) != 0;
    return v;
}

// exclusive scan of in[0..N) into out[0..N], out[N] = total (also stored to *total2 if given)
static int device_scan(Rank *c, const int *in, int N, int *out, int *total2) {
    int ntiles = (N + kScanTile - 1) / kScanTile;
    PS_LAUNCH(c, K_SCAN, 12.0 * N, "k_scan", {
        k_scan_tile_sums<<<ntiles, kThreads, 0, c->stream>>>(in, nullptr, N, c->tile_sums);
        k_scan_top<<<1, kThreads, 0, c->stream>>>(c->tile_sums, ntiles, out + N, total2);
        k_scan_apply<<<ntiles, kThreads, 0, c->stream>>>(in, nullptr, N, c->tile_sums, out);
    });
    if (c->opt_stats) c->stats.kernel_launches[K_SCAN] += 2;  // three kernels per scan
    return 0;
}

// both scans of the end of a step in three launches (k_scan2_*; two on grids of few tiles): a -> out_a (total also to *total_a2),
// acc + add -> out_b with acc += add, add = 0
static int device_scan2(Rank *c, const int *a, int *acc, int *add, int N, int *out_a, int *total_a2, int *out_b) {
    int ntiles = (N + kScanTile - 1) / kScanTile;
    if (ntiles <= kScanFewTiles && scan_few_ok()) {
        PS_LAUNCH(c, K_SCAN, 32.0 * N, "k_scan2_few", {
            k_scan2_tile_sums<<<ntiles, kThreads, 0, c->stream>>>(a, acc, add, N, ntiles, c->tile_sums);
            k_scan2_apply_few<<<ntiles, kThreads, 0, c->stream>>>(a, acc, add, N, ntiles, c->tile_sums, out_a, total_a2, out_b);
        });
        if (c->opt_stats) c->stats.kernel_launches[K_SCAN] += 1;
        return 0;
    }
    PS_LAUNCH(c, K_SCAN, 32.0 * N, "k_scan2", {
        k_scan2_tile_sums<<<ntiles, kThreads, 0, c->stream>>>(a, acc, add, N, ntiles, c->tile_sums);
        k_scan2_top<<<1, kThreads, 0, c->stream>>>(c->tile_sums, ntiles, out_a + N, total_a2, out_b + N);
        k_scan2_apply<<<ntiles, kThreads, 0, c->stream>>>(a, acc, add, N, ntiles, c->tile_sums, out_a, out_b);
    });
    if (c->opt_stats) c->stats.kernel_launches[K_SCAN] += 2;
    return 0;
}

// One stable radix pass over `bits` key bits starting at bit 0 of keyA (digits of <= 8 bits).
// Entries are (keyA, srcA); srcA == nullptr means source = index.  Results end in (keyA, srcA).
static int sort_passes(Rank *c, int n, int bits, unsigned *&keyA, unsigned *&keyB, int *&srcA, int *&srcB,
                       bool src_implicit) {
    const int passes = (bits + 7) / 8;
    const int dbits = (bits + passes - 1) / passes;
    const int ntiles = (n + kSortTile - 1) / kSortTile;
    for (int p = 0; p < passes; p++) {
        const int shift = p * dbits;
        const unsigned mask = (1u << dbits) - 1u;
        const int nb = (int)mask + 1;
        PS_LAUNCH(c, K_SORT_PASS, 4.0 * n, "k_sort_hist",
                  k_sort_hist<<<ntiles, kSortThreads, 0, c->stream>>>(n, keyA, shift, mask, ntiles, c->sort_hist));
        if (device_scan(c, c->sort_hist, nb * ntiles, c->sort_hist_scan, nullptr)) return 1;
        PS_LAUNCH(c, K_SORT_PASS, 16.0 * n, "k_sort_scatter",
                  k_sort_scatter<<<ntiles, kSortThreads, 0, c->stream>>>(
                      n, keyA, src_implicit ? nullptr : srcA, shift, mask, ntiles, c->sort_hist_scan, keyB, srcB));
        std::swap(keyA, keyB);
        std::swap(srcA, srcB);
        src_implicit = false;
    }
    return 0;
}

static int bits_for(unsigned long long max_value) {
    int b = 1;
    while (b < 63 && (max_value >> b)) b++;
    return b;
}

// Step 0: bring the uploaded particles (arbitrary slots) into (cell, id) order -- the binning of
// updateCOM (serial/parsim.cpp:265-285) done once; later steps keep the order by merging.
static int sort_initial(Rank *c, int *where = nullptr) {  // where (optional): slot -> ordered position
    const Geom &g = c->g;
    const int n = c->n_alive;
    if (rank_settle_rec_vel(c)) return 1;  // an upload with late velocities may still owe the records theirs
    if (n == 0) {
        PS_CUDA(c, cudaMemsetAsync(c->cs_cur, 0, sizeof(int) * ((size_t)g.ncells + 1), c->stream));
        return 0;
    }
    unsigned *keyA = (unsigned *)c->newcell, *keyB = (unsigned *)c->partner;
    int *srcA = c->arr_next, *srcB = c->mv_next;
    const int cell_bits = bits_for((unsigned long long)std::max(g.ncells - 1, 1));
    if (c->slots_id_ordered) {
        PS_LAUNCH(c, K_BIN_KEYS, 36.0 * n, "k_sort_cell_of",
                  k_sort_cell_of<<<grid_for(n), kThreads, 0, c->stream>>>(g, n, c->rec0, keyA));
        if (sort_passes(c, n, cell_bits, keyA, keyB, srcA, srcB, true)) return 1;
    } else {
        // slots are not in id order: sort by id first (LSD), then by cell
        PS_LAUNCH(c, K_BIN_KEYS, 36.0 * n, "k_sort_cell_of",
                  k_sort_cell_of<<<grid_for(n), kThreads, 0, c->stream>>>(g, n, c->rec0, c->tmp_id));
        PS_LAUNCH(c, K_SORT_PASS, 68.0 * n, "k_sort_id_keys",
                  k_sort_id_keys<<<grid_for(n), kThreads, 0, c->stream>>>(n, nullptr, c->rec0, keyA));
        if (sort_passes(c, n, bits_for((unsigned long long)std::max<long long>(c->n_part - 1, 1)), keyA, keyB,
                        srcA, srcB, true))
            return 1;
        PS_LAUNCH(c, K_SORT_PASS, 12.0 * n, "k_sort_cell_keys",
                  k_sort_cell_keys<<<grid_for(n), kThreads, 0, c->stream>>>(n, srcA, c->tmp_id, keyA));
        if (sort_passes(c, n, cell_bits, keyA, keyB, srcA, srcB, false)) return 1;
    }
    PS_LAUNCH(c, K_SORT_GATHER, 120.0 * n + 4.0 * g.ncells, "k_sort_gather",
              k_sort_gather<<<grid_for(n), kThreads, 0, c->stream>>>(
                  n, g.ncells, keyA, srcA, c->rec0, c->pos_alt, c->vel_alt, c->m_alt, c->id_alt, c->cell_alt,
                  c->cs_alt, where));
    std::swap(c->pos_cur, c->pos_alt);
    std::swap(c->vel_cur, c->vel_alt);
    std::swap(c->m_cur, c->m_alt);
    std::swap(c->id_cur, c->id_alt);
    std::swap(c->cell_cur, c->cell_alt);
    std::swap(c->cs_cur, c->cs_alt);
    return 0;
}

// ---- binning under the upload's shadow (host-buffer callers) ----------------------------------
// A single strip that receives its particles piece by piece in id order (parsim_upload) sorts every
// piece into (cell, id) order as soon as it is ingested -- on the step stream, while the copy engine
// brings the next piece over PCIe (6 ms per 8 Mi particles against 0.7 ms of sorting) -- and the
// first timestep only MERGES the pieces: per cell, the pieces' segments one after the other.  Same
// layout as sort_initial, bit for bit (tests/test_gpu_parity.py::test_upload_pieces_match_full_sort);
// parsim_reset and device-side init keep the full sort of the records.
static int sort_piece(Rank *c, int first, int cnt, int piece) {
    if (cnt <= 0) return 0;
    const Geom &g = c->g;
    unsigned *keyA = (unsigned *)c->newcell + first, *keyB = (unsigned *)c->partner + first;
    int *srcA = c->arr_next + first, *srcB = c->mv_next + first;
    const int cell_bits = bits_for((unsigned long long)std::max(g.ncells - 1, 1));
    PS_LAUNCH(c, K_BIN_KEYS, 36.0 * cnt, "k_sort_cell_of",
              k_sort_cell_of<<<grid_for(cnt), kThreads, 0, c->stream>>>(g, cnt, c->rec0 + first, keyA));
    if (sort_passes(c, cnt, cell_bits, keyA, keyB, srcA, srcB, true)) return 1;
    PS_LAUNCH(c, K_SORT_GATHER, 120.0 * cnt + 4.0 * g.ncells, "k_sort_gather",
              k_sort_gather<<<grid_for(cnt), kThreads, 0, c->stream>>>(
                  cnt, g.ncells, keyA, srcA, c->rec0 + first, c->pos_alt + first, c->vel_alt + first, c->m_alt + first,
                  c->id_alt + first, c->cell_alt + first, c->cs_pieces + (size_t)piece * ((size_t)g.ncells + 1)));
    return 0;
}

static int merge_pieces(Rank *c) {
    const Geom &g = c->g;
    const int P = (int)c->piece_first.size();
    if (P == 1) {  // one piece: it IS the sorted state
        PS_CUDA(c, cudaMemcpyAsync(c->cs_alt, c->cs_pieces, sizeof(int) * ((size_t)g.ncells + 1), cudaMemcpyDeviceToDevice,
                                   c->stream));
        std::swap(c->cs_cur, c->cs_alt);
    } else {
        PS_CUDA(c, cudaMemcpyAsync(c->d_piece_first, c->piece_first.data(), sizeof(int) * P, cudaMemcpyHostToDevice, c->stream));
        PS_LAUNCH(c, K_SORT_PASS, 8.0 * P * g.ncells, "k_piece_counts",
                  k_piece_counts<<<grid_for(g.ncells), kThreads, 0, c->stream>>>(g.ncells, P, c->cs_pieces, c->count));
        if (device_scan(c, c->count, g.ncells, c->cs_alt, nullptr)) return 1;
        PS_CUDA(c, cudaMemsetAsync(c->count, 0, sizeof(int) * (size_t)g.ncells, c->stream));  // the step loop expects zeros
        const double occ = (double)c->n_alive / std::max(1, g.ncells) / P;
        static const bool preload = !getenv("PARSIM_MERGE_PRELOAD") || atoi(getenv("PARSIM_MERGE_PRELOAD")) != 0;  // A/B
#define PS_MERGE_ARGS                                                                                                \
    g.ncells, P, c->d_piece_first, c->cs_pieces, c->cs_alt, c->pos_alt, c->vel_alt, c->m_alt, c->id_alt, c->pos_cur, \
        c->vel_cur, c->m_cur, c->id_cur, c->cell_cur
#define PS_MERGE(L)                                                                                          \
    do {                                                                                                     \
        if (P <= 16 && preload)                                                                              \
            k_merge_pieces<L, 16><<<grid_for((long long)g.ncells * L), kThreads, 0, c->stream>>>(PS_MERGE_ARGS); \
        else                                                                                                 \
            k_merge_pieces<L, 0><<<grid_for((long long)g.ncells * L), kThreads, 0, c->stream>>>(PS_MERGE_ARGS);  \
    } while (0)
        PS_LAUNCH(c, K_SORT_GATHER, 92.0 * c->n_alive + 8.0 * P * g.ncells, "k_merge_pieces", {
            if (occ <= 6) PS_MERGE(4); else if (occ <= 12) PS_MERGE(8); else if (occ <= 24) PS_MERGE(16); else PS_MERGE(32);
        });
#undef PS_MERGE
#undef PS_MERGE_ARGS
        std::swap(c->cs_cur, c->cs_alt);
        return 0;  // the merged state is already in the *_cur arrays
    }
    std::swap(c->pos_cur, c->pos_alt);
    std::swap(c->vel_cur, c->vel_alt);
    std::swap(c->m_cur, c->m_alt);
    std::swap(c->id_cur, c->id_alt);
    std::swap(c->cell_cur, c->cell_alt);
    return 0;
}

// step-0 binning: merge of the pre-sorted upload pieces when there are any, else the full sort
static int bin_initial(Rank *c) {
    if (c->piece_sorted && !c->piece_first.empty()) {
        c->piece_sorted = false;  // the piece arrays are consumed (they are the step loop's scratch)
        return merge_pieces(c);
    }
    return sort_initial(c);
}

template <int LANES>
static void launch_move(Rank *c, const MoveArgs &a, bool big, bool com) {
    long long threads = (long long)c->g.ncells * LANES;
    if (big) k_move<LANES, true, false><<<grid_for(threads), kThreads, 0, c->stream>>>(c->g, a);
    else if (com) k_move<LANES, false, true><<<grid_for(threads), kThreads, 0, c->stream>>>(c->g, a);
    else k_move<LANES, false, false><<<grid_for(threads), kThreads, 0, c->stream>>>(c->g, a);
}

// End of step: new cell_start (and tombstone_start) by scan, then the per-cell merge.
// Sources are entries [0, n_src) of (pos_alt, vel_cur, m_cur, id_cur) with their fate byte.
static int merge_step(Rank *c, int n_src, int n_surv, bool with_dead) {
    const Geom &g = c->g;
    if (with_dead && g.ncells <= 4000000) {
        // new cell_start and new tombstone_start (running tombstone counts += this step's) in ONE
        // three-launch scan: fewer launches where the grid is small enough for launches to matter
        // (at 1.25e7 cells the fused scan is 0.07 ms per step SLOWER than the two plain ones)
        if (device_scan2(c, c->count, c->dead_count, c->dead_new_count, g.ncells, c->cs_alt, &c->d_ctr->n_next,
                         c->dead_start_alt))
            return 1;
    } else if (with_dead) {
        if (device_scan(c, c->count, g.ncells, c->cs_alt, &c->d_ctr->n_next)) return 1;
        PS_LAUNCH(c, K_TOMBSTONES, 12.0 * g.ncells, "k_add_counts",
                  k_add_counts<<<grid_for(g.ncells), kThreads, 0, c->stream>>>(g.ncells, c->dead_count,
                                                                               c->dead_new_count));
        if (device_scan(c, c->dead_count, g.ncells, c->dead_start_alt, nullptr)) return 1;
    } else if (device_scan(c, c->count, g.ncells, c->cs_alt, &c->d_ctr->n_next)) {
        return 1;
    }
    MoveArgs a;
    a.cs_old = c->cs_cur; a.cs_new = c->cs_alt;
    a.ds_old = c->dead_start_cur; a.ds_new = with_dead ? c->dead_start_alt : nullptr;
    a.fate = c->fate;
    a.pos_in = c->pos_alt; a.vel_in = c->vel_cur; a.m_in = c->m_cur; a.id_in = c->id_cur;
    a.mv_head = c->mv_head; a.dm_head = c->dm_head; a.arr_head = c->arr_head; a.arr_cnt = c->arr_cnt; a.count = c->count;
    a.mv_next = c->mv_next;
    a.dead_old = c->dead_id_cur; a.dead_new = c->dead_id_alt;
    a.pos_out = c->pos_cur; a.vel_out = c->vel_alt; a.m_out = c->m_alt; a.id_out = c->id_alt;
    a.cell_out = c->cell_alt;
    a.big_list = c->move_big_list; a.big_count = c->d_move_big_count; a.big_min = c->move_big_min;
    a.com_xy = c->com_xy + g.nc; a.com_m = c->com_m + g.nc;
    // lanes per cell from the CURRENT occupancy (the layout does not depend on it)
    const double occ = (double)n_src / std::max(1, g.ncells);
    const int lanes = occ <= 6 ? 4 : occ <= 12 ? 8 : occ <= 24 ? 16 : 32;
    // huge cells / tombstone lists on average (BASELINE config 5): those go to k_move_big
    const bool big = c->move_big_force || occ > 0.125 * c->move_big_min ||
                     (with_dead && (double)c->n_tomb > 0.125 * c->move_big_min * g.ncells);
    // the next step's centres of mass come out of the same kernel when its lane groups are the
    // ones k_com would use (the lane count fixes the summation order); otherwise k_com runs
    const bool com = c->fuse_com && !big && lanes == c->com_lanes;
    // 1 fate + 4 id read per source entry, 40 read + 48 written per survivor, 8 B per tombstone
    PS_LAUNCH(c, K_MOVE, 5.0 * n_src + 88.0 * n_surv + (com ? 40.0 : 16.0) * g.ncells + (with_dead ? 8.0 * c->n_tomb : 0.0), "k_move", {
        switch (lanes) {
        case 4: launch_move<4>(c, a, big, com); break;
        case 8: launch_move<8>(c, a, big, com); break;
        case 16: launch_move<16>(c, a, big, com); break;
        default: launch_move<32>(c, a, big, com); break;
        }
        if (big) k_move_big<<<c->sm_count * 2, kThreads, 0, c->stream>>>(g, a);
    });
    c->com_valid = com;
    std::swap(c->vel_cur, c->vel_alt);
    std::swap(c->m_cur, c->m_alt);
    std::swap(c->id_cur, c->id_alt);
    std::swap(c->cell_cur, c->cell_alt);
    std::swap(c->cs_cur, c->cs_alt);
    if (with_dead) {
        std::swap(c->dead_start_cur, c->dead_start_alt);
        std::swap(c->dead_id_cur, c->dead_id_alt);
    }
    return 0;
}

template <int LANES>
static void launch_com(Rank *c) {
    long long threads = ((long long)c->g.ncells + kComCells - 1) / kComCells * LANES;
    k_com<LANES><<<grid_for(threads), kThreads, 0, c->stream>>>(
        c->g, c->pos_cur, c->m_cur, c->cs_cur, c->com_xy, c->com_m,
        c->opt_stats ? &c->d_ctr->force_pairs : nullptr);
}

// lanes per cell from the GLOBAL initial density (the same on every strip: the lane count fixes the
// summation order).  Letting the working lanes follow the current occupancy with the order kept
// (virtual lanes) was built and measured in round 2: 3.57 against 3.50 ms per config-3 run with 16
// lanes, no gain from the automatic choice (profiles/gpurun_r02_v8.log) -- removed again.
static void launch_com_any(Rank *c) {
    switch (c->com_lanes) {
    case 4: launch_com<4>(c); break;
    case 8: launch_com<8>(c); break;
    case 16: launch_com<16>(c); break;
    default: launch_com<32>(c); break;
    }
}

// Several strips: the particle count the last merge left (see the end of stage_collide).
static int settle_count(Rank *c) {
    if (!c->count_pending) return 0;
    c->count_pending = false;
    PS_CUDA(c, cudaEventSynchronize(c->ev_ring[0]));
    c->n_alive = c->h_ring[0].n_next;
    return 0;
}

// stage 0: binning (only when the cell order is not valid yet) + centre of mass + ghost rows
static int stage_bin_com(Rank *c) {
    const Geom &g = c->g;
    // with the centres of mass already there, nothing before the force stage needs the new count
    if (!(c->com_valid && c->sorted && !c->opt_stats) && settle_count(c)) return 1;
    if (!c->sorted) {
        if (bin_initial(c)) return 1;
        c->sorted = true;
    }
    // work-list counters of this step (a single strip's k_ghost_wrap below does it on the way)
    if (c->nranks > 1) k_step_reset<<<1, 1, 0, c->stream>>>(c->d_ctr, c->d_big_count, c->d_move_big_count);
    if (!c->com_valid) {
        PS_LAUNCH(c, K_COM, 24.0 * c->n_alive + 24.0 * g.ncells, "k_com", launch_com_any(c));
    } else if (c->opt_stats) {  // k_com also counts the ordered pairs of the force phase
        k_count_pairs<<<grid_for(g.ncells), kThreads, 0, c->stream>>>(g.ncells, c->cs_cur, &c->d_ctr->force_pairs);
        if (check_launch(c, "k_count_pairs")) return 1;
    }
    c->com_valid = false;  // consumed: the force phase moves the particles
    if (c->nranks == 1) {
        PS_LAUNCH(c, K_GHOST, 96.0 * g.nc, "k_ghost_wrap",
                  k_ghost_wrap<<<grid_for(g.nc), kThreads, 0, c->stream>>>(g, c->com_xy, c->com_m, c->d_ctr, c->d_big_count,
                                                                           c->d_move_big_count));
    } else if (!c->halo_valid && exchange_halo(c)) {  // (a collective strip upload has exchanged them already)
        return 1;
    }
    c->halo_valid = false;
    c->pos_latest = c->pos_cur;
    c->have_newcell = c->have_flags = false;
    return settle_count(c);
}

// stage 1: forces + integration
// the force sums alone (accelerations into `acc`): the first half of a step whose velocities are
// still on their way from the host (rank_upload_chunk); stage_force then only integrates
static int force_acc_only(Rank *c, double2 *acc) {
#define PS_FORCE_ACC(UN, TILE, CELLS, MINB, EARLY)                                                          \
    k_force_integrate<128, UN, TILE, CELLS, MINB, EARLY, true><<<grid_for(c->n_alive, 128), 128, 0, c->stream>>>( \
        c->g, c->n_alive, nullptr, c->pos_cur, c->m_cur, c->cell_cur, c->cs_cur, c->com_xy, c->com_m, nullptr,    \
        nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, acc)
    const int v = (double)c->n_alive < 28.0 * c->g.ncells ? c->sparse_variant : 2;  // as stage_force
    if (c->n_alive > 0)
    PS_LAUNCH(c, K_FORCE, 49.0 * c->n_alive + 24.0 * c->g.ncells, "k_force_integrate<acc>", {
        switch (v) {
        case 1: PS_FORCE_ACC(2, 256, 0, 16, false); break;
        case 5: PS_FORCE_ACC(2, 256, 0, 16, true); break;
        default: PS_FORCE_ACC(4, 768, 16, 8, false); break;
        }
    });
#undef PS_FORCE_ACC
    return 0;
}

static int stage_force(Rank *c) {
#define PS_FORCE_TB(TB, UN, TILE, CELLS, MINB, ...)                                                 \
    k_force_integrate<TB, UN, TILE, CELLS, MINB, ##__VA_ARGS__><<<grid_for(c->n_alive, TB), TB, 0, c->stream>>>( \
        c->g, c->n_alive, c->deferred ? &c->d_ctr->n_next : nullptr, c->pos_cur, c->m_cur, c->cell_cur, c->cs_cur, c->com_xy, c->com_m, \
        c->pos_alt, c->posf, c->vel_cur, c->newcell, c->flags, c->partner, &c->d_ctr->n_movers,     \
        c->dbg_acc)
#define PS_FORCE(UN, TILE, CELLS, MINB) PS_FORCE_TB(128, UN, TILE, CELLS, MINB)
    // variant by current occupancy (cells fill up or empty out as the run goes on)
    int v = c->force_variant;
    if (v <= 0) v = (double)c->n_alive < 28.0 * c->g.ncells ? c->sparse_variant : 2;  // measured crossover: about 30 per cell
    // an empty strip (everybody dead, or a strip the cloud does not reach) has nothing to move; the
    // kernels read cell[first] / cell[last] of their block unconditionally, so they are not launched
    if (c->acc0) {
        // the force sums of this step were taken while the velocities were uploaded
        double2 *acc = c->acc0;
        c->acc0 = nullptr;
        if (c->n_alive > 0)
        PS_LAUNCH(c, K_FORCE, 77.0 * c->n_alive, "k_integrate_acc",
                  k_integrate_acc<<<grid_for(c->n_alive), kThreads, 0, c->stream>>>(
                      c->g, c->n_alive, c->pos_cur, c->cell_cur, acc, c->pos_alt, c->posf, c->vel_cur, c->newcell,
                      c->flags, c->partner, &c->d_ctr->n_movers));
    } else if (c->n_alive > 0)
    PS_LAUNCH(c, K_FORCE, 81.0 * c->n_alive + 24.0 * c->g.ncells, "k_force_integrate", {
        switch (v) {
        case 1: PS_FORCE(2, 256, 0, 16); break;   // sparse cells: neighbourhood through L1, 32 registers, full occupancy
        case 3: PS_FORCE(8, 1280, 16, 5); break;  // deep unroll, 96 registers (comparison)
        case 4: PS_FORCE(2, 256, 0, 8); break;    // sparse cells at 64 registers (comparison)
        case 5: PS_FORCE_TB(128, 2, 256, 0, 16, true); break;  // sparse cells, neighbours before the barrier
        case 6: PS_FORCE_TB(128, 2, 256, 0, 12, true); break;  // the same at 40 registers
        case 10: PS_FORCE_TB(64, 4, 384, 8, 16); break;   // 2-warp blocks, 16 per SM: 29.7 ms per config-3 run (default 28.8)
        case 11: PS_FORCE_TB(256, 4, 1024, 16, 4); break; // 8-warp blocks, 4 per SM: 28.9 ms
        default: PS_FORCE(4, 768, 16, 8); break;  // 64 registers, 8 blocks per SM
        }
    });
#undef PS_FORCE
#undef PS_FORCE_TB
    c->pos_latest = c->pos_alt;
    c->have_newcell = true;
    return 0;
}

// Upload with late velocities (rank_upload_chunk): everything of the first timestep that needs
// positions and masses only -- the merge of the sorted pieces, the centres of mass, the force sums --
// runs while the velocity arrays are still crossing PCIe.  stage_bin_com / stage_force of the first
// step then find the centres of mass and the accelerations in place.
static int upload_prestep(Rank *c) {
    const Geom &g = c->g;
    if (bin_initial(c)) return 1;
    c->sorted = true;
    PS_LAUNCH(c, K_COM, 24.0 * c->n_alive + 24.0 * g.ncells, "k_com", launch_com_any(c));
    PS_LAUNCH(c, K_GHOST, 96.0 * g.nc, "k_ghost_wrap",
              k_ghost_wrap<<<grid_for(g.nc), kThreads, 0, c->stream>>>(g, c->com_xy, c->com_m, nullptr, nullptr, nullptr));
    c->com_valid = true;
    if (force_acc_only(c, c->vel_alt)) return 1;  // vel_alt is free until the merge at the end of the step
    c->acc0 = c->vel_alt;
    c->pos_latest = c->pos_cur;
    return 0;
}

// parsim_upload_strip on a multi-rank context (COLLECTIVE): all strips agree that their uploads went
// well, then every strip bins its particles, takes its centres of mass, exchanges the ghost rows and --
// when its velocities travel last (`late`) -- takes the force sums of the first timestep, all of it
// while the velocity arrays are still crossing PCIe.  *where_out: slot -> ordered position.
static int upload_prestep_strip(Rank *c, bool late, int **where_out) {
    const Geom &g = c->g;
    if (exchange_final_check(c)) return 1;
    int *where = late ? c->cell_cur : nullptr;  // the buffer that is cell_alt (free until the merge) after the sort's swap
    if (sort_initial(c, where)) return 1;
    c->sorted = true;
    *where_out = where;
    PS_LAUNCH(c, K_COM, 24.0 * c->n_alive + 24.0 * g.ncells, "k_com", launch_com_any(c));
    if (exchange_halo(c)) return 1;
    c->com_valid = true;
    c->halo_valid = true;
    if (late) {
        if (force_acc_only(c, c->vel_alt)) return 1;
        c->acc0 = c->vel_alt;
    }
    c->pos_latest = c->pos_cur;
    return 0;
}

// stage 2: list update (movers / strays), and emigrants to the neighbouring strips
static int stage_lists(Rank *c) {
    const Geom &g = c->g;
    // the mover list (written by k_force_integrate) lives in `partner`, which is free until the
    // collision search; about 1-5 % of the particles are on it
    if (c->nranks > 1) PS_CUDA(c, cudaMemsetAsync(c->mig_hdr_send, 0, 4 * sizeof(int), c->stream));
    PS_LAUNCH(c, K_CLASSIFY, 0.5 * c->n_alive, "k_classify",
              k_classify<<<c->sm_count * 4, kThreads, 0, c->stream>>>(
                  g, c->partner, &c->d_ctr->n_movers, c->cell_cur, c->cs_cur, c->newcell, c->id_cur,
                  c->n_tomb > 0 || c->deferred ? c->dead_start_cur : nullptr, c->dead_id_cur, c->flags, c->arr_head,
                  c->arr_next, c->pos_alt, c->arr_cnt, c->arr_rec, &c->d_ctr->strays, c->mig_list,
                  c->nranks > 1 ? c->mig_hdr_send + 3 : nullptr, c->mig_list_cap, c->d_ctr));
    c->n_guest = 0;
    if (c->nranks > 1 && exchange_emigrants(c)) return 1;
    c->have_flags = true;
    return 0;
}

template <int IT>
static void launch_collide_cells(Rank *c, const CollideArgs &a) {
    const size_t sh = kCollideWarps * warp_staging_bytes(32 * IT + 32);
    k_collide_cells<IT><<<grid_for(c->g.ncells, kCollideWarps), kCollideWarps * 32, sh, c->stream>>>(a);
}

// stage 3: collisions, compaction, and the (cell, id) order for the next step
static int stage_collide(Rank *c) {
    const Geom &g = c->g;
    const int n_tot = c->n_alive + c->n_guest;
    if (c->collide_grid) {
        CollideArgs a;
        a.g = g; a.n_res_val = c->n_alive; a.n_res_ptr = c->deferred ? &c->d_ctr->n_next : nullptr; a.thr2 = c->thr2; a.pos = c->pos_alt;
        a.cell = c->cell_cur; a.cell_start = c->cs_cur; a.newcell = c->newcell; a.id = c->id_cur;
        a.flags = c->flags; a.arr_head = c->arr_head; a.arr_next = c->arr_next; a.guest_src = c->guest_src;
        a.arr_cnt = c->arr_cnt; a.arr_rec = c->arr_rec;
        a.partner = c->partner;
        a.collision_pairs = c->opt_stats ? &c->d_ctr->collision_pairs : nullptr;
        a.worklist = c->big_cells; a.work_count = c->d_big_count;
        a.nb = c->grid_nb; a.inv_h = c->grid_nb / g.cellw; a.cap = c->grid_cap;
        PS_LAUNCH(c, K_COLLIDE, 29.0 * n_tot + 8.0 * g.ncells, "k_collide_cells", {
            // entries per lane from the CURRENT occupancy (+2 sigma); fuller cells go to k_collide_big
            const double now = (double)n_tot / std::max(1, g.ncells);
            const double want = now + 2.0 * sqrt(now) + 4.0;
            const int it = want <= 32 ? 1 : want <= 64 ? 2 : want <= 96 ? 3 : want <= 128 ? 4 : want <= 192 ? 6 : want <= 256 ? 8 : 12;
            switch (std::min(it, c->grid_it)) {
            case 1: launch_collide_cells<1>(c, a); break;
            case 2: launch_collide_cells<2>(c, a); break;
            case 3: launch_collide_cells<3>(c, a); break;
            case 4: launch_collide_cells<4>(c, a); break;
            case 6: launch_collide_cells<6>(c, a); break;
            case 8: launch_collide_cells<8>(c, a); break;
            default: launch_collide_cells<12>(c, a); break;
            }
        });
        a.nb = c->big_nb; a.inv_h = c->big_nb / g.cellw; a.cap = c->big_cap;
        PS_LAUNCH(c, K_COLLIDE_BIG, 0.0, "k_collide_big",
                  k_collide_big<<<c->sm_count, kThreads, staging_bytes(a.nb * a.nb, a.cap), c->stream>>>(
                      a, c->huge_items, c->d_big_count + 1));
        PS_LAUNCH(c, K_COLLIDE_BIG, 0.0, "k_collide_huge",
                  k_collide_huge<<<c->sm_count * 8, kThreads, 0, c->stream>>>(a, c->huge_items, c->d_big_count + 1));
    } else {
        PS_LAUNCH(c, K_COLLIDE, 21.0 * n_tot + 8.0 * g.ncells, "k_collide_find",
                  k_collide_find<<<grid_for(n_tot), kThreads, 0, c->stream>>>(
                      g, c->n_alive, c->deferred ? &c->d_ctr->n_next : nullptr, c->n_guest, c->thr2, c->thrf, c->pos_alt, c->posf, c->cell_cur, c->cs_cur, c->newcell,
                      c->id_cur, c->flags, c->arr_head, c->arr_next, c->guest_src, c->partner,
                      c->opt_stats ? &c->d_ctr->collision_pairs : nullptr));
    }
    if (c->nranks > 1) PS_CUDA(c, cudaMemsetAsync(c->mig_hdr_send, 0, 4 * sizeof(int), c->stream));
    PS_LAUNCH(c, K_RESOLVE, 10.0 * n_tot, "k_resolve",
              k_resolve<<<std::min(grid_for(n_tot, kThreads * kResolveRows), c->sm_count * 8), kThreads, 0, c->stream>>>(
                  g, c->n_alive, c->deferred ? &c->d_ctr->n_next : nullptr, c->n_guest, c->partner, c->cell_cur, c->newcell, c->flags, c->pos_alt,
                  c->vel_cur, c->id_cur, c->fate, c->count, c->dead_new_count, c->mv_head, c->dm_head,
                  c->mv_next, c->grave_pos, c->grave_vel, c->grave_id, c->grave_cell, c->cap, c->d_ctr, c->mig_list,
                  c->nranks > 1 ? c->mig_hdr_send + 3 : nullptr, c->mig_list_cap));
    if (c->deferred) {
        // Single strip, no instrumentation: the host does not wait for the step.  The kernels read the
        // particle count from Counters::n_next, the grids above are sized for the host's (older, hence
        // larger: a single strip only loses particles) count, the tombstone lists are rebuilt every
        // step, and the device error flag is looked at when the run ends (rank_run).
        if (merge_step(c, n_tot, n_tot, true)) return 1;
        Counters *slot = c->h_ring + (c->ring_pos % kCounterRing);
        PS_CUDA(c, cudaMemcpyAsync(slot, c->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
        if (!c->capturing) PS_CUDA(c, cudaEventRecord(c->ev_ring[c->ring_pos % kCounterRing], c->stream));
        c->ring_pos++;
        c->pos_latest = c->pos_cur;
        c->have_newcell = c->have_flags = false;
        c->steps_done++;
        c->stats.steps++;
        return 0;
    }
    int n_src = n_tot, tomb_in = 0;
    if (c->nranks > 1) {
        // several strips: the header round of the late exchange brings the counters to the host as well
        // (one host round trip instead of two), and a device error is not reported here but through
        // that exchange, which lets every strip leave together (parsim_exchange.cuh)
        if (exchange_late(c, &n_src, &tomb_in)) return 1;
    } else {
        PS_CUDA(c, cudaMemcpyAsync(c->h_ctr, c->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        if (c->h_ctr->error) return c->fail("device error code " + std::to_string(c->h_ctr->error));
    }
    const int grave_first = c->n_grave;
    const int deaths = c->h_ctr->grave_count - grave_first;
    c->n_grave = c->h_ctr->grave_count;
    // the tombstone lists change only when somebody died or a tombstone arrived: otherwise they
    // stay where they are (no scan, no copy)
    const bool with_dead = deaths > 0 || tomb_in > 0;
    if ((long long)c->n_tomb + deaths + tomb_in > c->cap) {
        const char *what = "strip capacity exceeded by tombstones (dead particles that lie in this strip): raise "
                           "PARSIM_CAPACITY_PCT";
        if (c->nranks == 1) return c->fail(what);
        exchange_soft_fail(c, what);
        c->n_tomb = 0;  // the tombstone lists are garbage from here on; the run ends at the next exchange
    }
    if (merge_step(c, n_src, n_src - deaths, with_dead)) return 1;
    c->n_tomb += deaths + tomb_in;  // upper bound; exact total is dead_start[ncells]
    c->stats.alive_particle_steps += c->n_alive;
    c->stats.movers += c->h_ctr->n_movers;
    c->stats.deaths += deaths;
    if (c->nranks == 1) {
        c->n_alive = n_tot - deaths;
    } else {
        // the strip's new particle count is the total of the scan: it is fetched without waiting, the next
        // step's ghost-row exchange is enqueued behind it, and the host picks it up before the force stage
        PS_CUDA(c, cudaMemcpyAsync(&c->h_ring[0].n_next, &c->d_ctr->n_next, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        PS_CUDA(c, cudaEventRecord(c->ev_ring[0], c->stream));
        c->count_pending = true;
    }
    c->n_guest = 0;
    c->pos_latest = c->pos_cur;
    c->have_newcell = c->have_flags = false;
    c->steps_done++;
    c->stats.steps++;
    return 0;
}

// dumps / downloads / results asked for before the first step: bring the uploaded records into
// the ordered layout now (stage 0 then skips its sort)
int rank_ensure_sorted(Rank *c) {
    PS_CUDA(c, cudaSetDevice(c->device));
    if (c->sorted || c->uploading) return 0;
    if (bin_initial(c)) return 1;
    c->sorted = true;
    c->pos_latest = c->pos_cur;
    return 0;
}

int rank_stage(Rank *c, int stage) {
    PS_CUDA(c, cudaSetDevice(c->device));
    if (c->uploading) return c->fail("upload not finished");
    if (stage != c->next_stage)
        return c->fail("stages must run in order; expected stage " + std::to_string(c->next_stage));
    int rc = 0;
    switch (stage) {
    case 0: rc = stage_bin_com(c); break;
    case 1: rc = stage_force(c); break;
    case 2: rc = stage_lists(c); break;
    case 3: rc = stage_collide(c); break;
    default: return c->fail("stage must be 0..3");
    }
    if (rc) return rc;
    c->next_stage = (stage + 1) & 3;
    return settle_count(c);  // callers between stages (dumps, results) see the final count
}

// The counters of the step before last, once they have arrived: the host stays at most two steps
// ahead of the device, and the occupancy-driven choices (kernel variants, lanes per cell) follow the
// real particle count with that lag.  A single strip never gains particles, so the older count is
// an upper bound for the grids.
static int deferred_refresh(Rank *c) {
    if (c->ring_pos < 2) return 0;
    const long long k = c->ring_pos - 2;
    PS_CUDA(c, cudaEventSynchronize(c->ev_ring[k % kCounterRing]));
    const Counters &h = c->h_ring[k % kCounterRing];
    if (h.error) return c->fail("device error code " + std::to_string(h.error));
    c->n_alive = h.n_next;
    c->n_grave = h.grave_count;
    c->n_tomb = h.grave_count;  // single strip: every dead particle lies in it
    return 0;
}

static int deferred_step_eager(Rank *c) {
    return stage_bin_com(c) || stage_force(c) || stage_lists(c) || stage_collide(c);
}

// What a replayed graph does NOT do: the host-side bookkeeping of the four stages (buffer roles
// after the merge, counters).  Must mirror merge_step and the deferred branch of stage_collide.
static void deferred_host_advance(Rank *c) {
    std::swap(c->vel_cur, c->vel_alt);
    std::swap(c->m_cur, c->m_alt);
    std::swap(c->id_cur, c->id_alt);
    std::swap(c->cell_cur, c->cell_alt);
    std::swap(c->cs_cur, c->cs_alt);
    std::swap(c->dead_start_cur, c->dead_start_alt);
    std::swap(c->dead_id_cur, c->dead_id_alt);
    c->n_guest = 0;
    c->acc0 = nullptr;  // consumed by the step's force stage, if it was there
    c->ring_pos++;
    c->pos_latest = c->pos_cur;
    c->have_newcell = c->have_flags = false;
    c->steps_done++;
    c->stats.steps++;
}

// One deferred step through a CUDA graph.  The graph bakes in pointers, grids and kernel choices:
// the key holds the roles of the ping-pong buffers, the counter slot and the occupancy class the
// choices are made from; the grids were sized for n_cap particles and stay valid while the bound on
// the particle count is not larger.  Whatever the key misses can only change WHICH equivalent kernel
// instantiation runs, never the result (tests/test_gpu_parity.py::test_force_kernel_variants_agree).
static int deferred_step_graph(Rank *c) {
    const double occ = (double)c->n_alive / std::max(1, c->g.ncells);
    const unsigned long long occ_class = (unsigned long long)std::min(63.0, std::max(0.0, 4.0 * log2(occ + 1.0)));
    const unsigned long long key = (c->pos_cur < c->pos_alt ? 1ull : 0) | (c->vel_cur < c->vel_alt ? 2ull : 0) |
                                   (c->dead_start_cur < c->dead_start_alt ? 4ull : 0) |
                                   ((unsigned long long)(c->ring_pos % kCounterRing) << 3) | (occ_class << 8) |
                                   (c->cs_cur < c->cs_alt ? 16ull : 0) |  // (the merge of upload pieces swaps the cell_start tables alone)
                                   (c->com_valid ? 1ull << 62 : 0) |  // does the step start with k_com?
                                   (c->acc0 ? 1ull << 61 : 0);        // ... and are its force sums already there?
    bool com_after = false;
    cudaGraphExec_t exec = nullptr;
    for (auto &sg : c->graphs)
        if (sg.key == key && sg.n_cap >= c->n_alive) { exec = sg.exec; com_after = sg.com_after; break; }
    if (!exec) {
        const int n_cap = c->n_alive;
        PS_CUDA(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
        c->capturing = true;
        const int rc = deferred_step_eager(c);  // records the launches, advances the host state
        c->capturing = false;
        cudaGraph_t graph = nullptr;
        const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
        if (rc) { if (graph) cudaGraphDestroy(graph); return 1; }
        if (e != cudaSuccess) return c->fail(std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e));
        PS_CUDA(c, cudaGraphInstantiate(&exec, graph, 0));
        cudaGraphDestroy(graph);
        if (c->graphs.size() >= 64) {  // bounded cache (long runs through many occupancy classes)
            cudaGraphExecDestroy(c->graphs.front().exec);
            c->graphs.erase(c->graphs.begin());
        }
        c->graphs.push_back({key, n_cap, exec, c->com_valid});
        PS_CUDA(c, cudaGraphLaunch(exec, c->stream));
        PS_CUDA(c, cudaEventRecord(c->ev_ring[(c->ring_pos - 1) % kCounterRing], c->stream));
        return 0;
    }
    PS_CUDA(c, cudaGraphLaunch(exec, c->stream));
    PS_CUDA(c, cudaEventRecord(c->ev_ring[c->ring_pos % kCounterRing], c->stream));
    deferred_host_advance(c);
    c->com_valid = com_after;  // what the replayed merge left behind
    return 0;
}

int rank_run(Rank *c, long n_steps) {
    PS_CUDA(c, cudaSetDevice(c->device));
    if (c->uploading) return c->fail("upload not finished");
    if (c->next_stage != 0) return c->fail("parsim_run in the middle of a debug-staged step");
    // No host round trip per step on a single strip whose steps are launch-bound (small problems),
    // unless per-launch statistics or per-launch error checks are on.  At 1e8 particles a step is
    // milliseconds of kernels, the round trip is 0.03 ms of it, and reading the particle count from
    // device memory costs the latency-bound kernels more than that (measured: 63.9 against 63.5 ms per
    // config-3 run, collision search +35 % at 8 per cell), so large problems keep the host in the loop.
    c->deferred = c->nranks == 1 && !c->opt_stats && !c->opt_sync_check && c->defer_steps;
    c->ring_pos = 0;
    PS_CUDA(c, cudaEventRecord(c->run_a, c->stream));
    for (long t = 0; t < n_steps; t++) {
        int rc;
        if (c->deferred) {
            rc = deferred_refresh(c);
            // the first step of a run may have to sort (eager); graphs only where a step is launch-bound
            if (!rc) rc = (c->use_graphs && c->sorted) ? deferred_step_graph(c) : deferred_step_eager(c);
        } else {
            rc = deferred_step_eager(c);
        }
        if (rc) { c->deferred = false; return 1; }
    }
    PS_CUDA(c, cudaEventRecord(c->run_b, c->stream));
    if (settle_count(c)) return 1;
    if (c->deferred && n_steps > 0) {
        PS_CUDA(c, cudaMemcpyAsync(c->h_ctr, c->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        c->deferred = false;
        if (c->h_ctr->error) return c->fail("device error code " + std::to_string(c->h_ctr->error));
        c->n_alive = c->h_ctr->n_next;
        c->n_grave = c->h_ctr->grave_count;
        c->n_tomb = c->h_ctr->grave_count;
    }
    c->deferred = false;
    if (c->nranks > 1 && n_steps > 0 && exchange_final_check(c)) return 1;
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0;
    PS_CUDA(c, cudaEventElapsedTime(&ms, c->run_a, c->run_b));
    c->stats.run_ms += ms;
    c->stats.particle_slots += (long long)c->n_part * n_steps;
    return timed_collect(c);
}

}  // namespace parsim

#include "parsim_dump.inl"
