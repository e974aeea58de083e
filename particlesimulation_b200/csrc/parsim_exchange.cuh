// parsim_exchange.cuh -- row-strip exchange over NCCL (NVLink 5 / NVSwitch).
// Re-expresses the reference's MPI decomposition (mpi/parsim-mpi.cpp:338-342 strips,
// :670-815 centre-of-mass ghost rows, :512-600 particle migration) with SERIAL semantics:
// the periodic wrap rows are really used (the MPI variant drops them), movers that
// the list update leaves behind are collision-checked on the strip they came from, and
// arrivals carry their global source cell so the destination list order is the serial one.
//
// Three exchange points per step, all neighbour-only (strip r <-> r-1, r+1, periodic):
//   halo      after the centre of mass: one cell row (3 doubles per cell) each way
//   emigrants after the list update:    removed movers whose new cell is in the other strip
//   late      after the collisions:     surviving strays that left the strip + tombstones
// Payloads are KB-scale; a record is 8 doubles.
// A close encounter (two particles a few 1e-6 apart) throws both many cells away in one step,
// exactly as in the reference; such "far" records (anything that is not the adjacent row of a
// neighbouring strip) are all-gathered and every strip keeps the ones that landed in its rows.
#pragma once
#include <limits.h>
#include <stdio.h>

#include "parsim_rank.cuh"
#include "parsim_merge.cuh"

namespace parsim {

constexpr int kFarCap = 4096;  // far records per strip per exchange (see exchange_records)
// One NCCL round per exchange: every message is a header record (the count) followed by a fixed
// "eager" number of record slots, so the receiver needs no count before it posts the receive.
// Counts above the eager size (rare: about 1e3 records cross a strip border per step at 1e8
// particles per GPU) travel in a second round that both sides know to post from the header.
constexpr int kEagerRecs = 8192;  // neighbour messages: 512 KB + header
constexpr int kFarEager = 32;     // all-gathered far records per strip in the first round
constexpr int kRec = 8;  // doubles per migrant record: x y vx vy m (id|src<<32) (cell|kind<<32) pad

__device__ __forceinline__ double pack_u64(unsigned long long v) { return __longlong_as_double((long long)v); }
__device__ __forceinline__ unsigned long long unpack_u64(double d) { return (unsigned long long)__double_as_longlong(d); }

// direction of a global cell that is outside this strip: 0 = up (row0-1), 1 = down, -1 = far
__device__ __forceinline__ int strip_dir(const Geom &g, int cell_global) {
    int row = cell_global / g.nc;
    if (row == (g.row0 - 1 + g.nc) % g.nc) return 0;
    if (row == (g.row0 + g.rows) % g.nc) return 1;
    return -1;
}

__device__ __forceinline__ void write_record(double *buf, int slot, double2 p, double2 v, double m,
                                             unsigned id, int src_cell, int cell, int kind) {
    double *r = buf + (size_t)slot * kRec;
    r[0] = p.x; r[1] = p.y; r[2] = v.x; r[3] = v.y; r[4] = m;
    r[5] = pack_u64((unsigned long long)id | ((unsigned long long)(unsigned)src_cell << 32));
    r[6] = pack_u64((unsigned long long)(unsigned)cell | ((unsigned long long)(unsigned)kind << 32));
    r[7] = 0.0;
}

// counts into the header record of the three send buffers
// The far header is all-gathered, so its second slot tells EVERY strip this strip's sticky device
// error code: all strips leave the step together instead of waiting for a strip that gave up.
__global__ void k_seal_headers(const int *__restrict__ hdr, double *send_up, double *send_dn, double *send_far,
                               const Counters *__restrict__ ctr) {
    if (threadIdx.x == 0) send_up[0] = pack_u64((unsigned long long)(unsigned)hdr[0]);
    if (threadIdx.x == 1) send_dn[0] = pack_u64((unsigned long long)(unsigned)hdr[1]);
    if (threadIdx.x == 2) {
        send_far[0] = pack_u64((unsigned long long)(unsigned)hdr[2]);
        send_far[1] = pack_u64((unsigned long long)(unsigned)ctr->error);
    }
}

// hdr[0] / hdr[1] = records for the strip above / below
// (over the short list k_classify wrote: hdr[3] entries)
__global__ void k_pack_emigrants(Geom g, const int *__restrict__ list, int list_cap, int mig_cap,
                                 const double2 *__restrict__ pos, const double2 *__restrict__ vel,
                                 const double *__restrict__ m, const unsigned *__restrict__ id,
                                 const int *__restrict__ cell, const int *__restrict__ newcell,
                                 double *send_up, double *send_dn, double *send_far, int *hdr,
                                 Counters *ctr) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= min(hdr[3], list_cap)) return;
    const int i = list[t];
    int dir = strip_dir(g, newcell[i]);
    if (dir < 0) dir = 2;  // far
    int slot = atomicAdd(&hdr[dir], 1);
    if (slot >= (dir == 2 ? kFarCap : mig_cap)) { ctr->error = DE_MIGRANT_OVERFLOW; return; }
    write_record(dir == 2 ? send_far : dir ? send_dn : send_up, slot, pos[i], vel[i], m[i], id[i],
                 cell[i] + g.cell_base, newcell[i], 0);
}

// guests: removed movers that arrived from the neighbouring strips; they sit behind the
// residents until the rebin and are linked into their destination cell's arrival bag.
__global__ void k_unpack_guests(Geom g, int n_rec, const double *__restrict__ recv, int first_idx,
                                int guest_first, double2 *pos, float2 *posf, double2 *vel, double *m, unsigned *id,
                                int *cell, int *newcell, unsigned char *flags, int *guest_src,
                                int *arr_head, int *arr_next, int *arr_cnt, int *partner, int filter,
                                Counters *ctr) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rec) return;
    const double *r = recv + (size_t)t * kRec;
    unsigned long long a = unpack_u64(r[5]), b = unpack_u64(r[6]);
    int idx = first_idx + t;
    int dstg = (int)(unsigned)(b & 0xffffffffu);
    int dst = dstg - g.cell_base;
    if (dst < 0 || dst >= g.ncells) {
        if (!filter) { ctr->error = 22; return; }
        // a far record that landed on another strip: a void entry that every kernel skips
        cell[idx] = -1;
        newcell[idx] = dstg;
        flags[idx] = F_MOVER | F_REMOVED | F_EMIGRANT | F_GUEST;
        partner[idx] = -1;
        return;
    }
    pos[idx] = make_double2(r[0], r[1]);
    if (posf) posf[idx] = make_float2((float)r[0], (float)r[1]);
    vel[idx] = make_double2(r[2], r[3]);
    m[idx] = r[4];
    id[idx] = (unsigned)(a & 0xffffffffu);
    cell[idx] = -1;
    newcell[idx] = dstg;
    flags[idx] = F_MOVER | F_REMOVED | F_GUEST;
    guest_src[guest_first + t] = (int)(unsigned)(a >> 32);
    arr_next[idx] = atomicExch(&arr_head[dst], idx);
    if (arr_cnt) atomicAdd(&arr_cnt[dst], 1 << 20);  // guests have no arrival record: the bag is walked
}

// after the collisions: strays that survived but now lie in another strip (kind 0) and
// tombstones of particles that died here but lie in another strip (kind 1)
// (over the short list k_resolve wrote: hdr[3] entries)
__global__ void k_pack_late(Geom g, const int *__restrict__ list, int list_cap, int mig_cap,
                            const unsigned char *__restrict__ fate, const double2 *__restrict__ pos,
                            const double2 *__restrict__ vel, const double *__restrict__ m,
                            const unsigned *__restrict__ id, const int *__restrict__ newcell,
                            double *send_up, double *send_dn, double *send_far, int *hdr, Counters *ctr) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= min(hdr[3], list_cap)) return;
    const int i = list[t];
    const int ft = fate[i];
    int dir = strip_dir(g, newcell[i]);
    if (dir < 0) dir = 2;  // far
    int slot = atomicAdd(&hdr[dir], 1);
    if (slot >= (dir == 2 ? kFarCap : mig_cap)) { ctr->error = DE_MIGRANT_OVERFLOW; return; }
    write_record(dir == 2 ? send_far : dir ? send_dn : send_up, slot, pos[i], vel[i], m[i], id[i], 0, newcell[i],
                 ft == FATE_LEFT_DEAD ? 1 : 0);
}

// late arrivals join the extended arrays behind residents and guests, as entries of the move
// bag (alive) or the dead bag (tombstone) of the cell they lie in
__global__ void k_unpack_late(Geom g, int n_rec, const double *__restrict__ recv, int first_idx,
                              double2 *pos, double2 *vel, double *m, unsigned *id, unsigned char *fate,
                              int *count, int *dead_new_count, int *mv_head, int *dm_head, int *mv_next,
                              int filter, Counters *ctr) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rec) return;
    const double *r = recv + (size_t)t * kRec;
    unsigned long long a = unpack_u64(r[5]), b = unpack_u64(r[6]);
    int idx = first_idx + t;
    int cg = (int)(unsigned)(b & 0xffffffffu);
    int kind = (int)(b >> 32);
    int c = cg - g.cell_base;
    if (c < 0 || c >= g.ncells) {
        if (!filter) { ctr->error = 24; return; }
        fate[idx] = FATE_EMIGRATED;  // landed on another strip: skipped by the merge
        return;
    }
    id[idx] = (unsigned)(a & 0xffffffffu);
    if (kind == 0) {
        pos[idx] = make_double2(r[0], r[1]);
        vel[idx] = make_double2(r[2], r[3]);
        m[idx] = r[4];
        fate[idx] = FATE_MOVE;
        atomicAdd(&count[c], 1);
        mv_next[idx] = atomicExch(&mv_head[c], idx);
    } else {
        fate[idx] = FATE_DEAD_MOVED;
        atomicAdd(&dead_new_count[c], 1);
        mv_next[idx] = atomicExch(&dm_head[c], idx);
    }
}

// ---- host side -----------------------------------------------------------------------------
#define PS_NCCL(c, call)                                                                   \
    do {                                                                                   \
        int r__ = (call);                                                                  \
        if (r__ != 0)                                                                      \
            return (c)->fail(std::string(#call) + ": " + nccl_api().GetErrorString(r__));  \
    } while (0)

inline int exchange_create(Rank *c, const unsigned char *nccl_id) {
    if (c->bus) {  // device-copy transport inside one process: no communicator
        PS_CUDA(c, cudaEventCreateWithFlags(&c->ev_bus, cudaEventDisableTiming));
    } else {
        NcclApi &api = nccl_api();
        if (!api.load()) return c->fail(api.error);
        if (!nccl_id) return c->fail("multi-rank context needs the NCCL unique id of rank 0");
        NcclUniqueId id;
        memcpy(id.internal, nccl_id, sizeof(id.internal));
        PS_NCCL(c, api.CommInitRank(&c->comm.comm, c->nranks, id, c->rank));
    }
    c->up = (c->rank - 1 + c->nranks) % c->nranks;
    c->dn = (c->rank + 1) % c->nranks;
    const size_t rec = ((size_t)std::max(c->mig_cap, kEagerRecs) + 1) * kRec;  // + header record
    for (int d = 0; d < 2; d++) {
        if (cudaMalloc((void **)&c->mig_send[d], rec * sizeof(double)) != cudaSuccess ||
            cudaMalloc((void **)&c->mig_recv[d], rec * sizeof(double)) != cudaSuccess)
            return c->fail("cudaMalloc of migrant buffers failed");
    }
    // hdr_send[0..2] = records for the strip above / below / far; hdr_recv[0..1]
    if (cudaMalloc((void **)&c->mig_hdr_send, 8 * sizeof(int)) != cudaSuccess)
        return c->fail("cudaMalloc failed");
    c->mig_hdr_recv = c->mig_hdr_send + 4;
    const size_t far = (size_t)kFarCap * kRec;
    if (cudaMalloc((void **)&c->far_send, (far + kRec) * sizeof(double)) != cudaSuccess ||
        cudaMalloc((void **)&c->far_recv, far * c->nranks * sizeof(double)) != cudaSuccess ||
        cudaMalloc((void **)&c->far_eager_recv, (size_t)(1 + kFarEager) * kRec * c->nranks * sizeof(double)) != cudaSuccess)
        return c->fail("cudaMalloc of far-record buffers failed");
    for (int d = 0; d < 2; d++) {  // slots beyond the count are sent too: keep them defined
        cudaMemset(c->mig_send[d], 0, rec * sizeof(double));
    }
    cudaMemset(c->far_send, 0, (far + kRec) * sizeof(double));
    if (cudaMalloc((void **)&c->guest_src, sizeof(int) * (2 * (size_t)c->mig_cap + (size_t)c->nranks * kFarCap)) != cudaSuccess)
        return c->fail("cudaMalloc failed");
    if (const char *e = getenv("PARSIM_EAGER_RECS")) {  // tests: force the overflow round (same value on every rank)
        c->eager_recs = std::min(std::max(atoi(e), 1), kEagerRecs);
        c->far_eager = std::min(c->eager_recs, kFarEager);
    }
    c->mig_list_cap = 2 * c->mig_cap + kFarCap;
    if (cudaMalloc((void **)&c->mig_list, sizeof(int) * (size_t)c->mig_list_cap) != cudaSuccess)
        return c->fail("cudaMalloc failed");
    if (cudaMallocHost((void **)&c->h_mig_hdr, (8 + c->nranks) * sizeof(int)) != cudaSuccess ||
        cudaMallocHost((void **)&c->h_hdr64, (2 + 2 * (size_t)c->nranks) * sizeof(unsigned long long)) != cudaSuccess)
        return c->fail("cudaMallocHost failed");
    return 0;
}

inline void exchange_destroy(Rank *c) {
    if (c->comm.comm) nccl_api().CommDestroy(c->comm.comm);
    c->comm.comm = nullptr;
    if (c->ev_bus) cudaEventDestroy(c->ev_bus);
    c->ev_bus = nullptr;
    for (int d = 0; d < 2; d++) {
        if (c->mig_send[d]) cudaFree(c->mig_send[d]);
        if (c->mig_recv[d]) cudaFree(c->mig_recv[d]);
    }
    if (c->mig_hdr_send) cudaFree(c->mig_hdr_send);
    if (c->far_send) cudaFree(c->far_send);
    if (c->far_recv) cudaFree(c->far_recv);
    if (c->far_eager_recv) cudaFree(c->far_eager_recv);
    if (c->mig_list) cudaFree(c->mig_list);
    if (c->h_mig_hdr) cudaFreeHost(c->h_mig_hdr);
    if (c->h_hdr64) cudaFreeHost(c->h_hdr64);
}

// ---- device-copy transport (LocalBus) ------------------------------------------------------
// Every strip PULLS what its neighbours would have sent: "ready" event on the own stream, host
// barrier, the own stream waits for the neighbours' events and copies out of their buffers (a
// plain device copy on one GPU, a peer copy over NVLink between GPUs).  The barrier after the
// copies are enqueued keeps a fast strip from re-recording its event before a slow one has
// started waiting for it.
static const char *const kPeerFailed = "another strip of the group failed";

inline int bus_copy(Rank *c, void *dst, const Rank *from, const void *src, size_t bytes) {
    if (!bytes) return 0;
    if (from->device == c->device)
        PS_CUDA(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, c->stream));
    else
        PS_CUDA(c, cudaMemcpyPeerAsync(dst, c->device, src, from->device, bytes, c->stream));
    return 0;
}

inline int bus_barrier(Rank *c) {
    if (c->bus->barrier()) return c->fail(kPeerFailed);
    return 0;
}

inline int bus_publish(Rank *c) {  // my buffers are final once the stream gets here
    PS_CUDA(c, cudaEventRecord(c->ev_bus, c->stream));
    return bus_barrier(c);
}

inline int local_halo(Rank *c) {
    const Geom &g = c->g;
    const size_t nc = (size_t)g.nc;
    Rank *up = c->bus->ranks[c->up], *dn = c->bus->ranks[c->dn];
    timed_begin(c, K_GHOST, 48.0 * nc);
    if (bus_publish(c)) return 1;
    PS_CUDA(c, cudaStreamWaitEvent(c->stream, up->ev_bus, 0));
    PS_CUDA(c, cudaStreamWaitEvent(c->stream, dn->ev_bus, 0));
    // my upper ghost row = the last row of the strip above, my lower one = the first row of the strip below
    if (bus_copy(c, c->com_xy, up, up->com_xy + (size_t)up->g.rows * nc, 2 * nc * sizeof(double)) ||
        bus_copy(c, c->com_m, up, up->com_m + (size_t)up->g.rows * nc, nc * sizeof(double)) ||
        bus_copy(c, c->com_xy + (g.rows + 1) * nc, dn, dn->com_xy + nc, 2 * nc * sizeof(double)) ||
        bus_copy(c, c->com_m + (g.rows + 1) * nc, dn, dn->com_m + nc, nc * sizeof(double)))
        return 1;
    timed_end(c);
    return bus_barrier(c);
}

// Ghost rows.  Receives are posted "from below" first so that with two strips (both
// neighbours are the same peer) sends and receives pair up in issue order: what I send up is
// the peer's row from below.
inline int exchange_halo(Rank *c) {
    if (c->bus) return local_halo(c);
    NcclApi &api = nccl_api();
    const Geom &g = c->g;
    const size_t nc = (size_t)g.nc;
    timed_begin(c, K_GHOST, 48.0 * nc);
    PS_NCCL(c, api.GroupStart());
    PS_NCCL(c, api.Send(c->com_xy + nc, 2 * nc, kNcclFloat64, c->up, c->comm.comm, c->stream));
    PS_NCCL(c, api.Send(c->com_m + nc, nc, kNcclFloat64, c->up, c->comm.comm, c->stream));
    PS_NCCL(c, api.Send(c->com_xy + g.rows * nc, 2 * nc, kNcclFloat64, c->dn, c->comm.comm, c->stream));
    PS_NCCL(c, api.Send(c->com_m + g.rows * nc, nc, kNcclFloat64, c->dn, c->comm.comm, c->stream));
    PS_NCCL(c, api.Recv(c->com_xy + (g.rows + 1) * nc, 2 * nc, kNcclFloat64, c->dn, c->comm.comm, c->stream));
    PS_NCCL(c, api.Recv(c->com_m + (g.rows + 1) * nc, nc, kNcclFloat64, c->dn, c->comm.comm, c->stream));
    PS_NCCL(c, api.Recv(c->com_xy, 2 * nc, kNcclFloat64, c->up, c->comm.comm, c->stream));
    PS_NCCL(c, api.Recv(c->com_m, nc, kNcclFloat64, c->up, c->comm.comm, c->stream));
    PS_NCCL(c, api.GroupEnd());
    timed_end(c);
    return 0;
}

// Error codes of all strips as they were when the headers were sealed (h_hdr64[2 + 2r + 1]).
inline int exchange_check_errors(Rank *c) {
    for (int r = 0; r < c->nranks; r++) {
        const int code = (int)c->h_hdr64[2 + 2 * r + 1];
        if (!code) continue;
        if (r == c->rank)
            return c->fail(!c->soft_err.empty() ? c->soft_err
                                                : "device error code " + std::to_string(code) +
                                                      " (1 = capacity, 3 = migrant buffer overflow; 22 / 24 = a record for another strip arrived)");
        return c->fail("strip " + std::to_string(r) + " reported device error code " + std::to_string(code));
    }
    return 0;
}

// A failure only THIS strip knows about (arrivals that do not fit): it becomes the strip's sticky
// device error, the strip carries on without the offending records, and the next exchange (or the
// check at the end of the run) tells everybody -- returning here would leave the other strips
// waiting in their next collective.
inline int exchange_soft_fail(Rank *c, const std::string &what) {
    if (c->soft_err.empty()) c->soft_err = what;
    k_set_int<<<1, 1, 0, c->stream>>>(&c->d_ctr->error, DE_CAPACITY);
    return 0;
}

// Ship hdr_send[0] records up, hdr_send[1] down and all-gather the hdr_send[2] far records, in ONE
// NCCL round (header + eager slots; see kEagerRecs).  On return mig_recv[0] / mig_recv[1] hold the
// header and then *from_up / *from_dn records from above / below, and far_base[r] points at the
// h_mig_hdr[8 + r] far records of strip r.
// h_mig_hdr: [0..2] sent up / down / far, [4..5] received from up / down, [8..8+nranks) far counts.
inline int exchange_records(Rank *c, int *from_up, int *from_dn, const double **far_base) {
    NcclApi &api = nccl_api();
    const size_t eager = (size_t)(1 + std::min(c->mig_cap, c->eager_recs)) * kRec;
    const size_t far_eager = (size_t)(1 + c->far_eager) * kRec;
    k_seal_headers<<<1, 32, 0, c->stream>>>(c->mig_hdr_send, c->mig_send[0], c->mig_send[1], c->far_send, c->d_ctr);
    PS_CUDA(c, cudaMemcpyAsync(c->h_mig_hdr, c->mig_hdr_send, 3 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    Rank *bup = c->bus ? c->bus->ranks[c->up] : nullptr, *bdn = c->bus ? c->bus->ranks[c->dn] : nullptr;
    if (c->bus) {
        // what the strip above sends DOWN is my message from above, and the other way round
        if (bus_publish(c)) return 1;
        for (Rank *q : c->bus->ranks)
            if (q != c) PS_CUDA(c, cudaStreamWaitEvent(c->stream, q->ev_bus, 0));
        if (bus_copy(c, c->mig_recv[0], bup, bup->mig_send[1], eager * sizeof(double)) ||
            bus_copy(c, c->mig_recv[1], bdn, bdn->mig_send[0], eager * sizeof(double)))
            return 1;
        for (Rank *q : c->bus->ranks)
            if (bus_copy(c, c->far_eager_recv + (size_t)q->rank * far_eager, q, q->far_send, far_eager * sizeof(double)))
                return 1;
    } else {
        PS_NCCL(c, api.GroupStart());
        PS_NCCL(c, api.Send(c->mig_send[0], eager, kNcclFloat64, c->up, c->comm.comm, c->stream));
        PS_NCCL(c, api.Send(c->mig_send[1], eager, kNcclFloat64, c->dn, c->comm.comm, c->stream));
        PS_NCCL(c, api.Recv(c->mig_recv[1], eager, kNcclFloat64, c->dn, c->comm.comm, c->stream));
        PS_NCCL(c, api.Recv(c->mig_recv[0], eager, kNcclFloat64, c->up, c->comm.comm, c->stream));
        PS_NCCL(c, api.AllGather(c->far_send, c->far_eager_recv, far_eager, kNcclFloat64, c->comm.comm, c->stream));
        PS_NCCL(c, api.GroupEnd());
    }
    PS_CUDA(c, cudaMemcpyAsync(c->h_hdr64 + 0, c->mig_recv[0], 8, cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaMemcpyAsync(c->h_hdr64 + 1, c->mig_recv[1], 8, cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaMemcpy2DAsync(c->h_hdr64 + 2, 16, c->far_eager_recv, far_eager * sizeof(double), 16, c->nranks,
                                 cudaMemcpyDeviceToHost, c->stream));  // (count, error code) of every strip
    PS_CUDA(c, cudaMemcpyAsync(c->h_ctr, c->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (exchange_check_errors(c)) return 1;  // the same verdict on every strip
    const int su = c->h_mig_hdr[0], sd = c->h_mig_hdr[1];
    const int ru = (int)c->h_hdr64[0], rd = (int)c->h_hdr64[1];
    c->h_mig_hdr[4] = ru;
    c->h_mig_hdr[5] = rd;
    if (ru > c->mig_cap || rd > c->mig_cap) return c->fail("migrant buffer overflow on receive");
    int far_max = 0;  // the same on every strip: everybody sees all the counts
    for (int r = 0; r < c->nranks; r++) {
        c->h_mig_hdr[8 + r] = (int)c->h_hdr64[2 + 2 * r];
        far_max = std::max(far_max, c->h_mig_hdr[8 + r]);
    }
    if (far_max > kFarCap) return c->fail("far-record buffer overflow");
    c->stats.migrants += su + sd + c->h_mig_hdr[2];
    // second round for what did not fit the eager slots (both ends know from the headers)
    const int E = (int)(eager / kRec) - 1;
    if (c->bus) {
        // the senders' buffers stay as they are until everybody has passed the barrier below, so
        // the rest is pulled without another handshake
        if (ru > E && bus_copy(c, c->mig_recv[0] + eager, bup, bup->mig_send[1] + eager, (size_t)(ru - E) * kRec * sizeof(double)))
            return 1;
        if (rd > E && bus_copy(c, c->mig_recv[1] + eager, bdn, bdn->mig_send[0] + eager, (size_t)(rd - E) * kRec * sizeof(double)))
            return 1;
        if (far_max > c->far_eager)
            for (Rank *q : c->bus->ranks)
                if (bus_copy(c, c->far_recv + (size_t)q->rank * far_max * kRec, q, q->far_send + kRec,
                             (size_t)far_max * kRec * sizeof(double)))
                    return 1;
        PS_CUDA(c, cudaStreamSynchronize(c->stream));  // my pulls are done: the senders may pack again
        if (bus_barrier(c)) return 1;
    } else if (su > E || sd > E || ru > E || rd > E || far_max > c->far_eager) {
        PS_NCCL(c, api.GroupStart());
        if (su > E) PS_NCCL(c, api.Send(c->mig_send[0] + eager, (size_t)(su - E) * kRec, kNcclFloat64, c->up, c->comm.comm, c->stream));
        if (sd > E) PS_NCCL(c, api.Send(c->mig_send[1] + eager, (size_t)(sd - E) * kRec, kNcclFloat64, c->dn, c->comm.comm, c->stream));
        if (rd > E) PS_NCCL(c, api.Recv(c->mig_recv[1] + eager, (size_t)(rd - E) * kRec, kNcclFloat64, c->dn, c->comm.comm, c->stream));
        if (ru > E) PS_NCCL(c, api.Recv(c->mig_recv[0] + eager, (size_t)(ru - E) * kRec, kNcclFloat64, c->up, c->comm.comm, c->stream));
        if (far_max > c->far_eager)
            PS_NCCL(c, api.AllGather(c->far_send + kRec, c->far_recv, (size_t)far_max * kRec, kNcclFloat64, c->comm.comm, c->stream));
        PS_NCCL(c, api.GroupEnd());
    }
    for (int r = 0; r < c->nranks; r++)
        far_base[r] = far_max > c->far_eager ? c->far_recv + (size_t)r * far_max * kRec
                                          : c->far_eager_recv + (size_t)r * far_eager + kRec;
    *from_up = ru;
    *from_dn = rd;
    return 0;
}

inline int exchange_emigrants(Rank *c) {
    const Geom &g = c->g;
    timed_begin(c, K_MIGRATE, 0);
    // hdr[3] = length of the emigrant list (k_classify); hdr[0..2] were zeroed with it
    k_pack_emigrants<<<grid_for(c->mig_list_cap), kThreads, 0, c->stream>>>(
        g, c->mig_list, c->mig_list_cap, c->mig_cap, c->pos_alt, c->vel_cur, c->m_cur, c->id_cur, c->cell_cur,
        c->newcell, c->mig_send[0] + kRec, c->mig_send[1] + kRec, c->far_send + kRec, c->mig_hdr_send, c->d_ctr);
    int ru = 0, rd = 0;
    std::vector<const double *> far_base(c->nranks);
    if (exchange_records(c, &ru, &rd, far_base.data())) return 1;
    int n_far = 0;
    for (int r = 0; r < c->nranks; r++) if (r != c->rank) n_far += c->h_mig_hdr[8 + r];
    if (c->n_alive + ru + rd + n_far > c->cap) {
        exchange_soft_fail(c, "strip capacity exceeded by arriving particles: raise PARSIM_CAPACITY_PCT");
        ru = rd = 0;
        for (int r = 0; r < c->nranks; r++) c->h_mig_hdr[8 + r] = 0;
        n_far = 0;
    }
    int first = c->n_alive, gfirst = 0;
    auto unpack = [&](int cnt, const double *buf, int filter) {
        if (!cnt) return;
        k_unpack_guests<<<grid_for(cnt), kThreads, 0, c->stream>>>(
            g, cnt, buf, first, gfirst, c->pos_alt, c->posf, c->vel_cur, c->m_cur, c->id_cur, c->cell_cur,
            c->newcell, c->flags, c->guest_src, c->arr_head, c->arr_next, c->arr_cnt, c->partner, filter, c->d_ctr);
        first += cnt;
        gfirst += cnt;
    };
    unpack(ru, c->mig_recv[0] + kRec, 0);
    unpack(rd, c->mig_recv[1] + kRec, 0);
    for (int r = 0; r < c->nranks; r++)
        if (r != c->rank) unpack(c->h_mig_hdr[8 + r], far_base[r], 1);
    timed_end(c);
    c->n_guest = ru + rd + n_far;
    return check_launch(c, "exchange_emigrants");
}

inline int exchange_late(Rank *c, int *n_src, int *n_tombstones_in) {
    const Geom &g = c->g;
    const int n_tot = c->n_alive + c->n_guest;
    timed_begin(c, K_MIGRATE, 0);
    // hdr[3] = length of the late list (k_resolve); hdr[0..2] were zeroed with it
    k_pack_late<<<grid_for(c->mig_list_cap), kThreads, 0, c->stream>>>(
        g, c->mig_list, c->mig_list_cap, c->mig_cap, c->fate, c->pos_alt, c->vel_cur, c->m_cur, c->id_cur, c->newcell,
        c->mig_send[0] + kRec, c->mig_send[1] + kRec, c->far_send + kRec, c->mig_hdr_send, c->d_ctr);
    int ru = 0, rd = 0;
    std::vector<const double *> far_base(c->nranks);
    if (exchange_records(c, &ru, &rd, far_base.data())) return 1;
    int n_far = 0;
    for (int r = 0; r < c->nranks; r++) if (r != c->rank) n_far += c->h_mig_hdr[8 + r];
    if (n_tot + ru + rd + n_far > c->cap) {
        exchange_soft_fail(c, "strip capacity exceeded by arriving particles: raise PARSIM_CAPACITY_PCT");
        ru = rd = 0;
        for (int r = 0; r < c->nranks; r++) c->h_mig_hdr[8 + r] = 0;
        n_far = 0;
    }
    int first = n_tot;
    auto unpack = [&](int cnt, const double *buf, int filter) {
        if (!cnt) return;
        k_unpack_late<<<grid_for(cnt), kThreads, 0, c->stream>>>(
            g, cnt, buf, first, c->pos_alt, c->vel_cur, c->m_cur, c->id_cur, c->fate, c->count,
            c->dead_new_count, c->mv_head, c->dm_head, c->mv_next, filter, c->d_ctr);
        first += cnt;
    };
    unpack(ru, c->mig_recv[0] + kRec, 0);
    unpack(rd, c->mig_recv[1] + kRec, 0);
    for (int r = 0; r < c->nranks; r++)
        if (r != c->rank) unpack(c->h_mig_hdr[8 + r], far_base[r], 1);
    timed_end(c);
    *n_src = first;
    *n_tombstones_in = ru + rd + n_far;  // upper bound
    return check_launch(c, "exchange_late");
}

// End of a run: do all strips agree that nothing went wrong?  (An error raised after the last
// exchange of the run -- by an unpack kernel, by the merge -- would otherwise surface on one strip
// only, and the others would wait for it in parsim_result's reduction.)
inline int exchange_final_check(Rank *c) {
    if (c->bus) {
        PS_CUDA(c, cudaMemcpyAsync(c->h_ctr, c->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        if (bus_barrier(c)) return 1;
        for (int r = 0; r < c->nranks; r++) c->h_hdr64[2 + 2 * r + 1] = (unsigned long long)(unsigned)c->bus->ranks[r]->h_ctr->error;
        if (bus_barrier(c)) return 1;
    } else {
        NcclApi &api = nccl_api();
        int *mine = c->mig_hdr_send + 4, *all = c->tile_sums;  // 1 int in, nranks ints out
        PS_CUDA(c, cudaMemcpyAsync(mine, &c->d_ctr->error, sizeof(int), cudaMemcpyDeviceToDevice, c->stream));
        PS_NCCL(c, api.AllGather(mine, all, 1, kNcclInt32, c->comm.comm, c->stream));
        std::vector<int> codes(c->nranks);
        PS_CUDA(c, cudaMemcpyAsync(codes.data(), all, sizeof(int) * c->nranks, cudaMemcpyDeviceToHost, c->stream));
        PS_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int r = 0; r < c->nranks; r++) c->h_hdr64[2 + 2 * r + 1] = (unsigned long long)(unsigned)codes[r];
    }
    return exchange_check_errors(c);
}

// sum of 4 doubles over all strips (particle 0 position, found flag, collisions)
inline int exchange_allreduce4(Rank *c, double v[4]) {
    NcclApi &api = nccl_api();
    double *d = c->mig_send[0];
    PS_CUDA(c, cudaMemcpyAsync(d, v, 4 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    PS_NCCL(c, api.AllReduce(d, d, 4, kNcclFloat64, kNcclSum, c->comm.comm, c->stream));
    PS_CUDA(c, cudaMemcpyAsync(v, d, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

}  // namespace parsim
