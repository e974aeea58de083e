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
    std::unique_ptr<LocalBus> bus;             // device-copy transport of the group (else NCCL)
    std::vector<cudaEvent_t> ev_ing;           // [strip][staging slot]: the strip has ingested the slot's piece
    long long group_pieces = 0;                // staging passes of the group upload so far
    bool shared_upload = false;                // group upload: ONE host->device copy per piece, every
                                               // strip ingests from it (same GPU or peer access)
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
        if (ctx->bus) ctx->bus->rearm();
        std::vector<std::thread> th;
        for (size_t r = 0; r < n; r++)
            th.emplace_back([&, r] {
                rc[r] = fn(ctx->ranks[r].get());
                if (rc[r] && ctx->bus) ctx->bus->abort();  // nobody waits for a strip that gave up
            });
        for (auto &t : th) t.join();
    }
    // report the strip that failed first-hand, not one that only saw the broken barrier
    size_t bad = n;
    for (size_t r = 0; r < n; r++)
        if (rc[r] && (bad == n || ctx->ranks[bad]->err == kPeerFailed)) bad = r;
    if (bad < n) return ctx_fail(ctx, (n > 1 ? "strip " + std::to_string(bad) + ": " : "") + ctx->ranks[bad]->err);
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
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        g_create_error = "no CUDA device: parsim_b200 has no CPU fallback";
        return 1;
    }
    // Transport of the group.  NCCL needs one GPU per strip; with fewer GPUs than strips (or
    // PARSIM_TRANSPORT=local) the strips share the GPUs round-robin and exchange by device copies
    // (LocalBus, parsim_rank.cuh) -- same kernels, same order rules, same bits.
    bool local = n_gpus > ndev;
    if (const char *t = getenv("PARSIM_TRANSPORT")) {
        if (!strcmp(t, "local")) local = true;
        else if (!strcmp(t, "nccl")) local = false;
        else { g_create_error = "PARSIM_TRANSPORT must be local or nccl"; return 1; }
    }
    if (!local && n_gpus > ndev) {
        g_create_error = "NCCL transport needs one GPU per strip (" + std::to_string(n_gpus) + " strips, " +
                         std::to_string(ndev) + " GPUs)";
        return 1;
    }
    unsigned char id[128];
    if (!local && parsim_nccl_unique_id(id)) return 1;
    std::unique_ptr<parsim_ctx> ctx(new parsim_ctx);
    ctx->group = true;
    for (int r = 0; r < n_gpus; r++) ctx->ranks.emplace_back(new Rank);
    if (local) {
        ctx->bus.reset(new LocalBus);
        for (auto &r : ctx->ranks) ctx->bus->ranks.push_back(r.get());
    }
    // peer access between the GPUs of the group: the group upload copies every piece to ONE GPU and
    // all strips ingest from there (NVLink reads), instead of n_gpus host->device copies of it
    const int used = std::min(n_gpus, ndev);
    ctx->shared_upload = true;
    for (int i = 0; i < used; i++)
        for (int j = 0; j < used; j++) {
            if (i == j) continue;
            int ok = 0;
            cudaDeviceCanAccessPeer(&ok, i, j);
            if (ok) {
                cudaSetDevice(i);
                cudaError_t e = cudaDeviceEnablePeerAccess(j, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = 0;
                cudaGetLastError();
            }
            if (!ok) ctx->shared_upload = false;
        }
    if (const char *e = getenv("PARSIM_SHARED_UPLOAD")) ctx->shared_upload = ctx->shared_upload && atoi(e) != 0;
    int rc = for_ranks(ctx.get(), [&](Rank *c) {
        int r = 0;
        while (ctx->ranks[r].get() != c) r++;
        return rank_create(c, side, ncside, n_part, r, n_gpus, r % ndev, local ? nullptr : id, first_rows, ctx->bus.get());
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
    for (cudaEvent_t e : ctx->ev_ing) if (e) cudaEventDestroy(e);
    delete ctx;
}

int parsim_upload_begin(parsim_ctx *ctx) { return for_ranks(ctx, rank_upload_begin); }

// Group upload: piece p is copied host->device ONCE, to the staging set of the first strip on GPU
// (p mod #GPUs) -- the GPUs' PCIe links work in parallel -- and every strip's ingest kernels read it
// from there (same GPU, or peer access over NVLink).  One host thread enqueues everything.
static int group_upload_chunk(parsim_ctx *ctx, long long id_base, const unsigned *ids, long long count,
                              const double *x, const double *y, const double *vx, const double *vy,
                              const double *m) {
    const size_t n = ctx->ranks.size();
    Rank *c0 = ctx->ranks[0].get();
    if (!c0->uploading) return ctx_fail(ctx, "parsim_upload_chunk outside upload_begin/upload_end");
    if (count < 0 || (!ids && (id_base < 0 || id_base + count > c0->n_part)))
        return ctx_fail(ctx, "upload chunk outside [0, n_part)");
    std::vector<Rank *> stagers;  // first strip on each GPU
    for (auto &r : ctx->ranks) {
        bool seen = false;
        for (Rank *q : stagers) seen = seen || q->device == r->device;
        if (!seen) stagers.push_back(r.get());
    }
    const size_t S = stagers.size();
    for (Rank *q : stagers)
        if (rank_stage_alloc(q)) return ctx_fail(ctx, q->err);
    if (ctx->ev_ing.empty()) {
        ctx->ev_ing.resize(n * 2 * S, nullptr);
        for (size_t r = 0; r < n; r++) {
            cudaSetDevice(ctx->ranks[r]->device);
            for (size_t k = 0; k < 2 * S; k++)
                if (cudaEventCreateWithFlags(&ctx->ev_ing[r * 2 * S + k], cudaEventDisableTiming) != cudaSuccess)
                    return ctx_fail(ctx, "cudaEventCreate failed");
        }
    }
    const double *src[5] = {x, y, vx, vy, m};
    const long long piece = stagers[0]->stage_cap;
#define PS_G(call)                                                                                         \
    do {                                                                                                   \
        cudaError_t e__ = (call);                                                                          \
        if (e__ != cudaSuccess) return ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)
    for (long long off = 0; off < count; off += piece) {
        const int cnt = (int)std::min<long long>(piece, count - off);
        const long long p = ctx->group_pieces;
        Rank *st = stagers[(size_t)(p % (long long)S)];
        const int b = (int)((p / (long long)S) & 1);
        const size_t slot = (size_t)(p % (long long)(2 * S));
        PS_G(cudaSetDevice(st->device));
        if (p >= (long long)(2 * S))  // the set is free once every strip has ingested its previous piece
            for (size_t r = 0; r < n; r++) PS_G(cudaStreamWaitEvent(st->copy_stream, ctx->ev_ing[r * 2 * S + slot], 0));
        for (int k = 0; k < 5; k++)
            PS_G(cudaMemcpyAsync(st->stage[b][k], src[k] + off, sizeof(double) * cnt, cudaMemcpyHostToDevice, st->copy_stream));
        if (ids)
            PS_G(cudaMemcpyAsync(st->stage_ids[b], ids + off, sizeof(unsigned) * cnt, cudaMemcpyHostToDevice, st->copy_stream));
        PS_G(cudaEventRecord(st->ev_copied[b], st->copy_stream));
        for (size_t r = 0; r < n; r++) {
            Rank *c = ctx->ranks[r].get();
            PS_G(cudaSetDevice(c->device));
            PS_G(cudaStreamWaitEvent(c->stream, st->ev_copied[b], 0));
            if (rank_ingest_piece(c, id_base + off, ids ? st->stage_ids[b] : nullptr, cnt, st->stage[b]))
                return ctx_fail(ctx, c->err);
            PS_G(cudaEventRecord(ctx->ev_ing[r * 2 * S + slot], c->stream));
        }
        ctx->group_pieces++;
    }
    for (Rank *q : stagers) {  // the caller's arrays are free again when this returns
        PS_G(cudaSetDevice(q->device));
        PS_G(cudaStreamSynchronize(q->copy_stream));
    }
#undef PS_G
    for (auto &r : ctx->ranks) r->n_uploaded += count;
    return 0;
}

int parsim_upload_chunk(parsim_ctx *ctx, long long id_base, long long count, const double *x,
                        const double *y, const double *vx, const double *vy, const double *m) {
    if (!x || !y || !vx || !vy || !m) return ctx_fail(ctx, "parsim_upload: NULL array");
    if (ctx->group && ctx->shared_upload) return group_upload_chunk(ctx, id_base, nullptr, count, x, y, vx, vy, m);
    return for_ranks(ctx, [&](Rank *c) { return rank_upload_chunk(c, id_base, nullptr, count, x, y, vx, vy, m); });
}

int parsim_upload_chunk_ids(parsim_ctx *ctx, long long count, const unsigned *ids, const double *x,
                            const double *y, const double *vx, const double *vy, const double *m) {
    if (!ids || !x || !y || !vx || !vy || !m) return ctx_fail(ctx, "parsim_upload_chunk_ids: NULL array");
    if (ctx->group && ctx->shared_upload) return group_upload_chunk(ctx, 0, ids, count, x, y, vx, vy, m);
    return for_ranks(ctx, [&](Rank *c) { return rank_upload_chunk(c, 0, ids, count, x, y, vx, vy, m); });
}

int parsim_upload_strip(parsim_ctx *ctx, long long count, const unsigned *ids, const double *x, const double *y,
                        const double *vx, const double *vy, const double *m) {
    if (ctx->ranks.size() != 1) return ctx_fail(ctx, "parsim_upload_strip: one strip per context only (one process per GPU)");
    Rank *c = ctx->ranks[0].get();  // (arguments are checked inside: a strip with bad ones must still reach the agreement)
    if (rank_upload_strip(c, count, ids, x, y, vx, vy, m)) return ctx_fail(ctx, c->err);
    return 0;
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
        // captured steps bake the options in (kernel arguments): drop them
        for (auto &sg : c->graphs) cudaGraphExecDestroy(sg.exec);
        c->graphs.clear();
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
