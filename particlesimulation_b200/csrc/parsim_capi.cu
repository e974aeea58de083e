// parsim_capi.cu -- the extern "C" boundary declared in include/parsim_b200.h.
// A parsim_ctx is either one strip of a multi-process job (parsim_create_rank) or a whole
// in-process job that owns n_gpus strips and drives them with one host thread each
// (parsim_create); n_gpus == 1 is simply one strip that spans the grid.
#include <string.h>

#include <functional>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "parsim_rank.cuh"
#include "parsim_exchange.cuh"
#include "parsim_host_init.hpp"

using namespace parsim;

struct parsim_ctx {
    std::vector<std::unique_ptr<Rank>> ranks;  // 1 for a rank context, n_gpus for a group
    bool group = false;                        // ranks are all the strips of the job
    std::string err;
};

static thread_local std::string g_create_error;

namespace {

int ctx_fail(parsim_ctx *ctx, const std::string &what) {
    ctx->err = what;
    return 1;
}

// run fn(rank) for every strip, concurrently when there are several (NCCL calls of the
// strips must progress together)
int for_ranks(parsim_ctx *ctx, const std::function<int(Rank *)> &fn) {
    const size_t n = ctx->ranks.size();
    std::vector<int> rc(n, 0);
    if (n == 1) {
        rc[0] = fn(ctx->ranks[0].get());
    } else {
        std::vector<std::thread> th;
        for (size_t r = 0; r < n; r++)
            th.emplace_back([&, r] { rc[r] = fn(ctx->ranks[r].get()); });
        for (auto &t : th) t.join();
    }
    for (size_t r = 0; r < n; r++)
        if (rc[r]) return ctx_fail(ctx, (n > 1 ? "gpu " + std::to_string(r) + ": " : "") + ctx->ranks[r]->err);
    return 0;
}

}  // namespace

extern "C" {

int parsim_abi_version(void) { return PARSIM_ABI_VERSION; }

const char *parsim_last_error(const parsim_ctx *ctx) {
    return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

int parsim_nccl_unique_id(unsigned char id_out[128]) {
    NcclApi &api = nccl_api();
    if (!api.load()) { g_create_error = api.error; return 1; }
    NcclUniqueId id;
    int r = api.GetUniqueId(&id);
    if (r != 0) { g_create_error = std::string("ncclGetUniqueId: ") + api.GetErrorString(r); return 1; }
    memcpy(id_out, id.internal, 128);
    return 0;
}

int parsim_create_rank_rows(parsim_ctx **out, double side, long ncside, long long n_part, int rank,
                            int nranks, int device, const unsigned char *nccl_unique_id,
                            const long *first_rows) {
    if (!out) { g_create_error = "out is NULL"; return 1; }
    *out = nullptr;
    std::unique_ptr<parsim_ctx> ctx(new parsim_ctx);
    ctx->ranks.emplace_back(new Rank);
    if (rank_create(ctx->ranks[0].get(), side, ncside, n_part, rank, nranks, device, nccl_unique_id, first_rows)) {
        g_create_error = ctx->ranks[0]->err;
        rank_destroy(ctx->ranks[0].get());
        return 1;
    }
    *out = ctx.release();
    return 0;
}

int parsim_create_rank(parsim_ctx **out, double side, long ncside, long long n_part, int rank,
                       int nranks, int device, const unsigned char *nccl_unique_id) {
    return parsim_create_rank_rows(out, side, ncside, n_part, rank, nranks, device, nccl_unique_id, nullptr);
}

int parsim_create_rows(parsim_ctx **out, double side, long ncside, long long n_part, int n_gpus,
                       const long *first_rows) {
    if (!out) { g_create_error = "out is NULL"; return 1; }
    *out = nullptr;
    if (n_gpus < 1) { g_create_error = "n_gpus must be >= 1"; return 1; }
    if (n_gpus == 1) return parsim_create_rank_rows(out, side, ncside, n_part, 0, 1, 0, nullptr, first_rows);
    unsigned char id[128];
    if (parsim_nccl_unique_id(id)) return 1;
    std::unique_ptr<parsim_ctx> ctx(new parsim_ctx);
    ctx->group = true;
    for (int r = 0; r < n_gpus; r++) ctx->ranks.emplace_back(new Rank);
    int rc = for_ranks(ctx.get(), [&](Rank *c) {
        int r = 0;
        while (ctx->ranks[r].get() != c) r++;
        return rank_create(c, side, ncside, n_part, r, n_gpus, r, id, first_rows);
    });
    if (rc) {
        g_create_error = ctx->err;
        for (auto &r : ctx->ranks) rank_destroy(r.get());
        return 1;
    }
    *out = ctx.release();
    return 0;
}

int parsim_create(parsim_ctx **out, double side, long ncside, long long n_part, int n_gpus) {
    return parsim_create_rows(out, side, ncside, n_part, n_gpus, nullptr);
}

int parsim_balance_rows(long ncside, int nranks, const long long *row_counts, long *first_rows) {
    if (ncside < 1 || nranks < 1 || nranks > ncside || !row_counts || !first_rows) {
        g_create_error = "parsim_balance_rows: bad arguments";
        return 1;
    }
    for (long r = 0; r < ncside; r++)
        if (row_counts[r] < 0) { g_create_error = "parsim_balance_rows: negative count"; return 1; }
    parsim::balance_rows(ncside, nranks, row_counts, first_rows);
    return 0;
}

int parsim_host_row_counts(double side, long ncside, long long count, const double *y, long long *row_counts) {
    if (!(side > 0) || ncside < 1 || count < 0 || !y || !row_counts) {
        g_create_error = "parsim_host_row_counts: bad arguments";
        return 1;
    }
    const double cellw = side / (double)ncside;  // serial/parsim.cpp:268-272
    for (long long i = 0; i < count; i++) {
        long r = (long)(int)(y[i] / cellw);
        r = r < 0 ? 0 : r >= ncside ? ncside - 1 : r;
        row_counts[r]++;
    }
    return 0;
}

void parsim_destroy(parsim_ctx *ctx) {
    if (!ctx) return;
    for (auto &r : ctx->ranks) rank_destroy(r.get());
    delete ctx;
}

int parsim_upload_begin(parsim_ctx *ctx) { return for_ranks(ctx, rank_upload_begin); }

int parsim_upload_chunk(parsim_ctx *ctx, long long id_base, long long count, const double *x,
                        const double *y, const double *vx, const double *vy, const double *m) {
    if (!x || !y || !vx || !vy || !m) return ctx_fail(ctx, "parsim_upload: NULL array");
    return for_ranks(ctx, [&](Rank *c) { return rank_upload_chunk(c, id_base, nullptr, count, x, y, vx, vy, m); });
}

int parsim_upload_chunk_ids(parsim_ctx *ctx, long long count, const unsigned *ids, const double *x,
                            const double *y, const double *vx, const double *vy, const double *m) {
    if (!ids || !x || !y || !vx || !vy || !m) return ctx_fail(ctx, "parsim_upload_chunk_ids: NULL array");
    return for_ranks(ctx, [&](Rank *c) { return rank_upload_chunk(c, 0, ids, count, x, y, vx, vy, m); });
}

int parsim_reset(parsim_ctx *ctx) { return for_ranks(ctx, rank_reset); }

int parsim_download(parsim_ctx *ctx, long long capacity, unsigned *ids, double *x, double *y, double *vx,
                    double *vy, double *m, long long *count) {
    if (ctx->ranks.size() != 1) return ctx_fail(ctx, "parsim_download: one strip per context only");
    if (!ids || !x || !y || !vx || !vy || !m || !count) return ctx_fail(ctx, "parsim_download: NULL array");
    Rank *c = ctx->ranks[0].get();
    if (rank_download(c, capacity, ids, x, y, vx, vy, m, count)) return ctx_fail(ctx, c->err);
    return 0;
}

int parsim_fp64_peak(parsim_ctx *ctx, double *tflops) {
    Rank *c = ctx->ranks[0].get();
    if (!tflops) return ctx_fail(ctx, "tflops is NULL");
    if (rank_fp64_peak(c, tflops)) return ctx_fail(ctx, c->err);
    return 0;
}

int parsim_pair_peak(parsim_ctx *ctx, double *pairs_per_s) {
    Rank *c = ctx->ranks[0].get();
    if (!pairs_per_s) return ctx_fail(ctx, "pairs_per_s is NULL");
    if (rank_pair_peak(c, pairs_per_s)) return ctx_fail(ctx, c->err);
    return 0;
}

int parsim_upload_end(parsim_ctx *ctx) { return for_ranks(ctx, rank_upload_end); }

int parsim_upload(parsim_ctx *ctx, const double *x, const double *y, const double *vx,
                  const double *vy, const double *m) {
    if (parsim_upload_begin(ctx)) return 1;
    if (parsim_upload_chunk(ctx, 0, ctx->ranks[0]->n_part, x, y, vx, vy, m)) return 1;
    return parsim_upload_end(ctx);
}

int parsim_host_init_particles(long seed, double side, long ncside, long long n_part, long long first,
                               long long count, double *x, double *y, double *vx, double *vy, double *m) {
    if (!x || !y || !vx || !vy || !m || first < 0 || count < 0 || first + count > n_part || ncside < 1) {
        g_create_error = "parsim_host_init_particles: bad arguments";
        return 1;
    }
    parsim_host::RandomStream rng((int)seed);  // main stores the seed in an int, :465
    for (long long i = 0; i < 5 * first; i++) rng.next();
    parsim_host::init_particles(rng, side, ncside, n_part, count, x, y, vx, vy, m);
    return 0;
}

int parsim_strip_rows(long ncside, int nranks, int rank, long *first_row, long *n_rows) {
    if (ncside < 1 || nranks < 1 || rank < 0 || rank >= nranks || nranks > ncside || !first_row || !n_rows) {
        g_create_error = "parsim_strip_rows: bad arguments";
        return 1;
    }
    parsim::strip_rows(ncside, nranks, rank, first_row, n_rows);
    return 0;
}

int parsim_init_device(parsim_ctx *ctx, long seed) {
    return for_ranks(ctx, [&](Rank *c) { return rank_init_device(c, seed); });
}

int parsim_run(parsim_ctx *ctx, long n_steps) {
    if (n_steps < 0) return ctx_fail(ctx, "n_steps must be >= 0");
    return for_ranks(ctx, [&](Rank *c) { return rank_run(c, n_steps); });
}

int parsim_result(parsim_ctx *ctx, double *x0, double *y0, long long *collisions) {
    const size_t n = ctx->ranks.size();
    std::vector<double> px(n), py(n);
    std::vector<int> found(n);
    std::vector<long long> coll(n);
    int rc = for_ranks(ctx, [&](Rank *c) {
        size_t r = 0;
        while (ctx->ranks[r].get() != c) r++;
        return rank_result_local(c, &px[r], &py[r], &found[r], &coll[r]);
    });
    if (rc) return rc;
    double v[4] = {0, 0, 0, 0};
    for (size_t r = 0; r < n; r++) {
        if (found[r]) { v[0] += px[r]; v[1] += py[r]; v[2] += 1; }
        v[3] += (double)coll[r];
    }
    Rank *c0 = ctx->ranks[0].get();
    if (!ctx->group && c0->nranks > 1) {
        if (exchange_allreduce4(c0, v)) return ctx_fail(ctx, c0->err);
    }
    if (v[2] != 1.0) return ctx_fail(ctx, "particle 0 found " + std::to_string((int)v[2]) + " times");
    if (x0) *x0 = v[0];
    if (y0) *y0 = v[1];
    if (collisions) *collisions = (long long)v[3];
    return 0;
}

int parsim_debug_stage(parsim_ctx *ctx, int stage) {
    return for_ranks(ctx, [&](Rank *c) { return rank_stage(c, stage); });
}

int parsim_dump(parsim_ctx *ctx, double *x, double *y, double *vx, double *vy, double *m, int *alive,
                int *cell, int *list_cell, unsigned long long *list_rank, unsigned char *owned) {
    // strips are dumped one after the other; each fills only the particles it owns
    const size_t n = (size_t)ctx->ranks[0]->n_part;
    if (owned) memset(owned, 0, n);
    for (auto &r : ctx->ranks) {
        std::vector<unsigned char> own(owned ? n : 0);
        if (rank_dump(r.get(), x, y, vx, vy, m, alive, cell, list_cell, list_rank, owned ? own.data() : nullptr))
            return ctx_fail(ctx, r->err);
        if (owned) for (size_t i = 0; i < n; i++) owned[i] |= own[i];
    }
    return 0;
}

int parsim_dump_cells(parsim_ctx *ctx, double *mx, double *my, double *m) {
    if (ctx->ranks.size() != 1) return ctx_fail(ctx, "parsim_dump_cells: single-strip contexts only");
    Rank *c = ctx->ranks[0].get();
    if (rank_dump_cells(c, mx, my, m)) return ctx_fail(ctx, c->err);
    return 0;
}

const char *parsim_kernel_name(int k) {
    static const char *names[K_COUNT_] = {
        "ingest", "bin_keys", "scan", "sort_passes", "sort_gather", "com", "ghost_rows",
        "force_integrate", "classify_movers", "collide_find", "collide_resolve", "tombstones",
        "migrate", "init_particles", "merge_move", "collide_big_cells"};
    return (k >= 0 && k < K_COUNT_) ? names[k] : nullptr;
}

int parsim_stats(parsim_ctx *ctx, parsim_stats_t *out) {
    if (!out) return ctx_fail(ctx, "out is NULL");
    memset(out, 0, sizeof(*out));
    for (auto &rp : ctx->ranks) {
        Rank *c = rp.get();
        const parsim_stats_t &s = c->stats;
        out->steps = s.steps;
        out->particle_slots = s.particle_slots;  // global n_part * steps on every strip
        out->alive_particle_steps += s.alive_particle_steps;
        out->force_pairs += (long long)c->h_ctr->force_pairs;
        out->collision_pairs += (long long)c->h_ctr->collision_pairs;
        out->movers += s.movers;
        out->strays += (long long)c->h_ctr->strays;
        out->deaths += s.deaths;
        out->migrants += s.migrants;
        out->run_ms = s.run_ms > out->run_ms ? s.run_ms : out->run_ms;
        for (int k = 0; k < PARSIM_NKERNELS; k++) {
            out->kernel_ms[k] = s.kernel_ms[k] > out->kernel_ms[k] ? s.kernel_ms[k] : out->kernel_ms[k];
            out->kernel_launches[k] += s.kernel_launches[k];
            out->kernel_bytes[k] += s.kernel_bytes[k];
        }
    }
    return 0;
}

int parsim_stats_reset(parsim_ctx *ctx) {
    return for_ranks(ctx, [&](Rank *c) {
        PS_CUDA(c, cudaSetDevice(c->device));
        memset(&c->stats, 0, sizeof(c->stats));
        PS_CUDA(c, cudaMemsetAsync(&c->d_ctr->force_pairs, 0, 4 * sizeof(unsigned long long), c->stream));
        c->h_ctr->force_pairs = c->h_ctr->collision_pairs = c->h_ctr->movers = c->h_ctr->strays = 0;
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    });
}

int parsim_set_option(parsim_ctx *ctx, const char *name, long long value) {
    if (!name) return ctx_fail(ctx, "option name is NULL");
    std::string n(name);
    for (auto &rp : ctx->ranks) {
        Rank *c = rp.get();
        if (n == "stats") c->opt_stats = value != 0;
        else if (n == "sync_check") c->opt_sync_check = value != 0;
        else if (n == "keep_initial") c->opt_keep_initial = value != 0;
        else if (n == "semantics") {
            if (value != 0 && value != 1) return ctx_fail(ctx, "semantics: 0 = serial (default), 1 = omp");
            if (c->next_stage != 0) return ctx_fail(ctx, "semantics cannot change in the middle of a step");
            c->g.no_strays = (int)value;
        }
        else if (n == "debug_acc") {
            if (value && !c->dbg_acc) {
                cudaSetDevice(c->device);
                if (cudaMalloc((void **)&c->dbg_acc, sizeof(double2) * (size_t)c->cap) != cudaSuccess)
                    return ctx_fail(ctx, "cudaMalloc(debug_acc) failed");
            }
        } else return ctx_fail(ctx, "unknown option: " + n);
    }
    return 0;
}

void *parsim_stream(parsim_ctx *ctx) { return (void *)ctx->ranks[0]->stream; }

}  // extern "C"
