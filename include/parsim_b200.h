/*
 * parsim_b200.h -- C ABI of the B200-native per-timestep loop of `parsim`.
 *
 * The reference (MiguelParece/ParticleSimulation) has no library/plugin/FFI interface: each
 * variant is one executable whose `main` constructs `ParticleSimulation`, calls
 * `simulate(n_step)` and prints the result (serial/parsim.cpp:457-489).  This header is the
 * boundary a maintainer would bind instead of the four phase functions called from
 * `ParticleSimulation::simulate` (serial/parsim.cpp:422-448); INTEGRATION.md shows the
 * patch.  Plain pointers and sizes only, no C++/torch types.  All functions return 0 on
 * success; on failure they return non-zero and parsim_last_error() gives the text that
 * `main` prints as "Error: ..." before exiting 1 (mirrors serial/parsim.cpp:484-488).
 * No exception crosses this boundary.  A context is driven by one host thread at a time.
 *
 * There is no CPU fallback: every entry point needs a CUDA device (sm_100a build).
 */
#ifndef PARSIM_B200_H
#define PARSIM_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct parsim_ctx parsim_ctx;

#define PARSIM_ABI_VERSION 1
int parsim_abi_version(void);

/* ---- life cycle -------------------------------------------------------------------------
 * Replaces the ParticleSimulation constructor's allocations (serial/parsim.cpp:209-216:
 * particles(n_part), cellParticles(ncside^2), cells(ncside^2)).  `n_gpus` > 1 splits the
 * grid into n_gpus row strips inside this process (one host thread per strip inside
 * parsim_run; mpi/parsim-mpi.cpp:338-342,:371-372 is the reference's equivalent
 * decomposition).  With at least n_gpus devices every strip gets its own GPU and the strips
 * exchange halos / migrants over NCCL; with fewer devices the strips share the GPUs
 * round-robin and exchange by device-to-device copies -- the reference's own way of testing
 * its distributed variant on one machine (mpi/run_tests.sh:9-17).  Results do not depend on
 * n_gpus or on the transport (environment PARSIM_TRANSPORT=local|nccl overrides the choice). */
int parsim_create(parsim_ctx **out, double side, long ncside, long long n_part, int n_gpus);

/* One strip of a multi-process job (one process per GPU, e.g. under torchrun): this process
 * is `rank` of `nranks`, uses CUDA device `device`, and joins the NCCL communicator named by
 * the 128-byte `nccl_unique_id` that rank 0 obtained from parsim_nccl_unique_id() and
 * shipped to the others out of band.  nranks == 1 needs no id (NULL). */
int parsim_nccl_unique_id(unsigned char id_out[128]);
int parsim_create_rank(parsim_ctx **out, double side, long ncside, long long n_part,
                       int rank, int nranks, int device, const unsigned char *nccl_unique_id);

/* Explicit strips (equal-PARTICLE strips for clustered inputs -- the case the
 * reference's report calls "Uneven"): strip r owns grid rows [first_rows[r], first_rows[r+1]),
 * first_rows[0] = 0, first_rows[nranks] = ncside, strictly increasing.  NULL = equal-row strips.
 * parsim_balance_rows derives such boundaries from per-row particle counts (prefix sum; every
 * strip keeps >= 1 row); parsim_host_row_counts ADDS the rows of `count` y coordinates to
 * row_counts[ncside] with the reference's binning arithmetic (serial/parsim.cpp:268-272).  The
 * last two are host arithmetic and need no GPU. */
int parsim_create_rows(parsim_ctx **out, double side, long ncside, long long n_part, int n_gpus,
                       const long *first_rows);
int parsim_create_rank_rows(parsim_ctx **out, double side, long ncside, long long n_part,
                            int rank, int nranks, int device, const unsigned char *nccl_unique_id,
                            const long *first_rows);
int parsim_balance_rows(long ncside, int nranks, const long long *row_counts, long *first_rows);
int parsim_host_row_counts(double side, long ncside, long long count, const double *y,
                           long long *row_counts);

void parsim_destroy(parsim_ctx *ctx);
const char *parsim_last_error(const parsim_ctx *ctx); /* ctx may be NULL: last create error */

/* ---- initial condition ------------------------------------------------------------------
 * Replaces the state that init_particles leaves in `particles` (serial/parsim.cpp:220-232).
 * Arrays are host memory, length n_part, particle id = array index; they are copied.
 * Every mass must be > 0 (true for every seed of the reference generator).  In a
 * multi-rank job every rank passes the same full arrays and keeps its own strip; an
 * in-process group (parsim_create, n_gpus > 1) copies every piece to ONE of its GPUs and all
 * strips ingest from there (peer access), i.e. the host->device volume does not grow with
 * the number of strips. */
int parsim_upload(parsim_ctx *ctx, const double *x, const double *y, const double *vx,
                  const double *vy, const double *m);
/* Same, in pieces: particles [id_base, id_base+count).  Call parsim_upload_begin first and
 * parsim_upload_end last; lets a host stream 8e8 particles without holding them all. */
int parsim_upload_begin(parsim_ctx *ctx);
int parsim_upload_chunk(parsim_ctx *ctx, long long id_base, long long count, const double *x,
                        const double *y, const double *vx, const double *vy, const double *m);
int parsim_upload_end(parsim_ctx *ctx);
/* Pieces with explicit particle ids (a rank of a multi-process job that already holds only
 * its own strip's particles in host memory); same begin/end bracket. */
int parsim_upload_chunk_ids(parsim_ctx *ctx, long long count, const unsigned *ids, const double *x,
                            const double *y, const double *vx, const double *vy, const double *m);
/* The WHOLE strip of this rank in one call (begin + chunk_ids + end), for one-process-per-GPU jobs
 * (what mpi/parsim-mpi.cpp:338-342 does with MPI_Scatterv of the initial particles, here from each
 * rank's own host memory).  On a context created with parsim_create_rank* and more than one rank the
 * call is COLLECTIVE -- every rank makes it, like parsim_run: the strips agree that all uploads went
 * well (one fails, all fail), bin their particles, take their centres of mass and exchange the
 * ghost rows before it returns, and x, y, m, ids travel before vx, vy so that this work (and the
 * force sums of the first timestep) runs while the velocities are still crossing PCIe.  The
 * arrays are free again when it returns.  One strip per context. */
int parsim_upload_strip(parsim_ctx *ctx, long long count, const unsigned *ids, const double *x,
                        const double *y, const double *vx, const double *vy, const double *m);
/* The alive particles of this context's strip, in device order, into host arrays of
 * `capacity` elements; *count = how many.  One strip per context. */
int parsim_download(parsim_ctx *ctx, long long capacity, unsigned *ids, double *x, double *y,
                    double *vx, double *vy, double *m, long long *count);
/* Restart from the state last uploaded (the uploaded records stay in device memory untouched
 * by the step loop): no host traffic.  The option "keep_initial" is accepted and ignored. */
int parsim_reset(parsim_ctx *ctx);
/* Host-side (C++) restatement of RandomGenerator + init_particles (serial/parsim.cpp:18-48,
 * :220-232): fills particles [first, first+count) of the seeded stream into host arrays of
 * `count` elements.  Needs no GPU.  Both distributions (seed < 0: truncated normal). */
int parsim_host_init_particles(long seed, double side, long ncside, long long n_part,
                               long long first, long long count, double *x, double *y,
                               double *vx, double *vy, double *m);
/* Row strips (mpi/parsim-mpi.cpp:338-342 splits the grid rows the same way): rank r of nranks
 * owns grid rows [first_row, first_row + n_rows).  Host arithmetic only, needs no GPU; what
 * parsim_create / parsim_create_rank use. */
int parsim_strip_rows(long ncside, int nranks, int rank, long *first_row, long *n_rows);
/* Device-side restatement of init_particles for seeds >= 0 (uniform stream; xorshift32
 * jump-ahead, bit-identical to serial/parsim.cpp:26-32,:220-232).  Negative seeds
 * (rejection sampling, :34-43) return an error: generate on the host and upload. */
int parsim_init_device(parsim_ctx *ctx, long seed);

/* ---- the hot path -----------------------------------------------------------------------
 * Replaces ParticleSimulation::simulate (serial/parsim.cpp:422-448): n_steps times
 * updateCOM, updateForces, updatePositionAndVelocity, checkCollisions.  Returns after the
 * device has finished (the reference times exactly this call, :475-477).  May be called
 * repeatedly; steps accumulate. */
int parsim_run(parsim_ctx *ctx, long n_steps);

/* Replaces print_result's reads (serial/parsim.cpp:450-453): position of particle 0
 * (frozen if it has collided) and the running collision total.  In a multi-rank job every
 * rank gets the global values. */
int parsim_result(parsim_ctx *ctx, double *x0, double *y0, long long *collisions);

/* ---- test / measurement hooks -----------------------------------------------------------
 * parsim_debug_stage runs ONE stage of the next step so tests can compare the device state
 * with the reference after the matching phase:
 *   stage 0 = binning + centre of mass              (updateCOM            :261-290)
 *   stage 1 = forces + integration + wrap + new cell (updateForces :292-377, applyForce :150-195)
 *   stage 2 = list update (movers, strays)           (updateCellParticles :234-259)
 *   stage 3 = collisions + compaction                (checkCollisions     :388-420)
 * Stages must be called in order 0,1,2,3,0,...  Single-rank contexts only. */
int parsim_debug_stage(parsim_ctx *ctx, int stage);

/* Id-ordered copy of the particle state (any pointer may be NULL; arrays of n_part).
 * alive[i] = 1/0; list_cell[i] = cell whose list particle i is in for the collision phase
 * (valid after stage 2; -1 for dead particles); list_rank[i] = position key inside that
 * list (ascending = list order).  Multi-rank: every rank fills only the particles it owns and
 * leaves the rest untouched; owned[i] says which. */
int parsim_dump(parsim_ctx *ctx, double *x, double *y, double *vx, double *vy, double *m,
                int *alive, int *cell, int *list_cell, unsigned long long *list_rank,
                unsigned char *owned);
/* Per-cell centre of mass after stage 0 (arrays of ncside^2, row-major like the reference's
 * `cells`; single-rank only). */
int parsim_dump_cells(parsim_ctx *ctx, double *mx, double *my, double *m);

/* Counters and per-kernel device times since the last parsim_stats_reset().  Kernel times are
 * CUDA-event measurements on the stream the kernels are launched on and are only collected
 * after parsim_set_option(ctx, "stats", 1). */
#define PARSIM_NKERNELS 16
typedef struct parsim_stats_t {
    long long steps;
    long long particle_slots;       /* n_part * steps (the metric's numerator)            */
    long long alive_particle_steps; /* sum over steps of particles alive at step start    */
    long long force_pairs;          /* ordered in-cell pair evaluations in the force kernel */
    long long collision_pairs;      /* ordered pair tests in the collision kernel          */
    long long movers;               /* particles that changed cell                         */
    long long strays;               /* movers left behind by the list-update rule          */
    long long deaths;
    long long migrants;             /* particles sent to another strip                     */
    double run_ms;                  /* device time of parsim_run calls (events)            */
    double kernel_ms[PARSIM_NKERNELS];
    long long kernel_launches[PARSIM_NKERNELS];
    double kernel_bytes[PARSIM_NKERNELS]; /* algorithmic bytes           */
} parsim_stats_t;
const char *parsim_kernel_name(int k); /* NULL past the last kernel */
int parsim_stats(parsim_ctx *ctx, parsim_stats_t *out);
int parsim_stats_reset(parsim_ctx *ctx);

/* FP64 FMA-chain microbenchmark on this context's GPU (the FP64 roofline denominator). */
int parsim_fp64_peak(parsim_ctx *ctx, double *tflops);
/* Pair interactions per second of the force loop's own instruction mix (13 FP64 + 1 MUFU per
 * pair) on register operands: the ceiling the force kernel's inner loop is compared against. */
int parsim_pair_peak(parsim_ctx *ctx, double *pairs_per_s);

/* options: "keep_initial" (0/1, see parsim_reset), "stats" (0/1), "sync_check" (0/1: synchronise +
 * check after every kernel), "semantics" (0 = the serial program, the default and the parity
 * target; 1 = the OpenMP variant's list update, openMP/parsim-omp.cpp:248-278, which moves every
 * leaver -- no strays -- and is otherwise identical; pinned against parsim-omp run with ONE
 * thread, whose list order is deterministic.  The MPI variant's predicate d^2 <= EPSILON2,
 * mpi/parsim-mpi.cpp:1015, is not offered: it cannot be built here, so it could not be pinned). */
int parsim_set_option(parsim_ctx *ctx, const char *name, long long value);

/* Raw device pointer of the CUDA stream the kernels run on (cudaStream_t), for callers that
 * want to record their own events. */
void *parsim_stream(parsim_ctx *ctx);

/* Centre of mass of every cell (the result of stage 0: cell masses and mass-weighted mean
 * positions, as parsim_debug_stage(ctx, 0) + parsim_dump_cells give them) WITHOUT ANY ALLOCATION,
 * for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it.  This entry point has no counterpart in the
 * reference.
 *   d_x, d_y, d_m      DEVICE pointers, n_part doubles each (particle i = index i, as parsim_upload)
 *   d_mx, d_my, d_mass DEVICE pointers, ncside * ncside doubles each (cell = row * ncside + column;
 *                      an empty cell gets 0, 0, 0)
 *   stream             cudaStream_t to launch on (NULL = the legacy default stream); the call is
 *                      stream-ordered and returns without synchronising
 * All pointers must belong to the CURRENT device.  The binning scratch is a fixed device array
 * inside the library (about 4 MiB, held by every GPU that loads the library until the process
 * exits), hence the limits: n_part <= PARSIM_STATIC_MAX_PARTICLES and ncside * ncside <=
 * PARSIM_STATIC_MAX_CELLS; anything larger is rejected (returns 1, text in parsim_last_error(NULL))
 * -- use a parsim_ctx there.  No cudaMalloc / cudaFree / allocating library routine is called.
 * Every call on a device shares the array: calls that overlap on the GPU (two streams, two
 * threads) race -- serialise them.  The first call on a device looks the array's address up
 * (cudaGetSymbolAddress); make one call (n_part = 0 will do) before starting a stream capture.
 * Results equal the context path's bit for bit (same per-cell order, same summation tree). */
#define PARSIM_STATIC_MAX_PARTICLES 262144
#define PARSIM_STATIC_MAX_CELLS 262144
int parsim_cell_com_static(double side, long ncside, long long n_part, const double *d_x,
                           const double *d_y, const double *d_m, double *d_mx, double *d_my,
                           double *d_mass, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PARSIM_B200_H */
