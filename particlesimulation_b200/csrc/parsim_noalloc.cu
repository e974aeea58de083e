// parsim_noalloc.cu -- parsim_cell_com_static: the centre-of-mass phase (binning + per-cell sums,
// the work of stage 0) for callers that may not allocate: code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc.  It has no counterpart in the
// reference.
//
// The context-based path (parsim_upload + parsim_debug_stage(0)) bins with the radix sort of
// parsim_sort.cuh, whose keys / sources / histograms live in cudaMalloc'ed buffers owned by the
// context.  This variant bins with a counting sort whose scratch is the file-scope array below:
// nothing is allocated, at set-up or per call.  Each cell's particles end up in index order (the
// reference's bin-time list order, serial/parsim.cpp:265-285) and are summed in the order k_com
// sums them, so the results are the context path's bit for bit.
//
// The array is shared by every call on a device and keeps the last call's contents; the entry
// point clears what must start at zero.  Calls that overlap on one GPU (different streams, or
// different host threads) race on it -- callers serialise them.
#include "parsim_kernels.cuh"
#include "../../include/parsim_b200.h"

namespace parsim {

constexpr int kStaticMaxParticles = PARSIM_STATIC_MAX_PARTICLES;
constexpr int kStaticMaxCells = PARSIM_STATIC_MAX_CELLS;

struct StaticScratch {
    int cell_start[kStaticMaxCells + 64];  // exclusive scan of the counts, [ncells] = n
    int cursor[kStaticMaxCells];           // counts, then the fill cursor of every cell
    int bag[kStaticMaxParticles];          // particle indices per cell in arrival (atomic) order
    int sorted[kStaticMaxParticles];       // the same lists in index order
};

// ~4 MiB, held by every GPU that loads the library until the process exits
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

__device__ __forceinline__ int static_cell_of(const Geom &g, const double *x, const double *y, int i) {
    return global_cell(make_double2(x[i], y[i]), g);
}

__global__ void k_static_clear(int ncells, int *cursor) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncells) cursor[c] = 0;
}

__global__ void k_static_count(Geom g, int n, const double *__restrict__ x, const double *__restrict__ y,
                               int *cursor) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicAdd(&cursor[static_cell_of(g, x, y, i)], 1);
}

// one block: cell_start = exclusive scan of the counts, cursor = cell_start
__global__ void __launch_bounds__(1024) k_static_scan(int ncells, int *cursor, int *cell_start) {
    __shared__ int part[1024];
    const int per = (ncells + 1023) / 1024;
    const int first = min(threadIdx.x * per, ncells), last = min(first + per, ncells);
    int sum = 0;
    for (int c = first; c < last; c++) sum += cursor[c];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    int run = part[threadIdx.x] - sum;
    for (int c = first; c < last; c++) {
        const int k = cursor[c];
        cell_start[c] = run;
        cursor[c] = run;
        run += k;
    }
    if (threadIdx.x == 1023) cell_start[ncells] = part[1023];
}

__global__ void k_static_fill(Geom g, int n, const double *__restrict__ x, const double *__restrict__ y,
                              int *cursor, int *__restrict__ bag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) bag[atomicAdd(&cursor[static_cell_of(g, x, y, i)], 1)] = i;
}

// rank of every entry inside its cell = how many of the cell's entries have a smaller index
__global__ void k_static_order(Geom g, int n, const double *__restrict__ x, const double *__restrict__ y,
                               const int *__restrict__ cell_start, const int *__restrict__ bag,
                               int *__restrict__ sorted) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int i = bag[k];
    const int c = static_cell_of(g, x, y, i);
    const int s = cell_start[c], e = cell_start[c + 1];
    int rank = 0;
    for (int j = s; j < e; j++) rank += bag[j] < i;
    sorted[s + rank] = i;
}

// k_com's arithmetic (parsim_kernels.cuh): LANES lanes per cell, lane l sums entries l, l + LANES,
// ... of the cell's list with the same fma chain, then the same xor-shuffle tree and division.
template <int LANES>
__global__ void __launch_bounds__(kThreads)
k_static_com(int ncells, const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ m,
             const int *__restrict__ cell_start, const int *__restrict__ sorted, double *__restrict__ mx,
             double *__restrict__ my, double *__restrict__ mass) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = t / LANES, lane = t % LANES;
    const int s = cell_start[min(c, ncells)], e = cell_start[min(c + 1, ncells)];
    double M = 0, Mx = 0, My = 0;
    for (int k = s + lane; k < e; k += LANES) {
        const int i = sorted[k];
        const double mv = m[i];
        M += mv;
        Mx = fma(mv, x[i], Mx);
        My = fma(mv, y[i], My);
    }
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) {
        M += __shfl_xor_sync(0xffffffffu, M, o);
        Mx += __shfl_xor_sync(0xffffffffu, Mx, o);
        My += __shfl_xor_sync(0xffffffffu, My, o);
    }
    if (lane == 0 && c < ncells) {
        mx[c] = M > 0 ? Mx / M : 0.0;
        my[c] = M > 0 ? My / M : 0.0;
        mass[c] = M;
    }
}

static int static_fail(const std::string &msg) {
    g_create_error = "parsim_cell_com_static: " + msg;
    return 1;
}

}  // namespace parsim

using namespace parsim;

extern "C" int parsim_cell_com_static(double side, long ncside, long long n_part, const double *d_x,
                                      const double *d_y, const double *d_m, double *d_mx, double *d_my,
                                      double *d_mass, void *stream) {
    if (!(side > 0)) return static_fail("side must be positive");
    if (ncside < 1 || (long long)ncside * ncside > kStaticMaxCells)
        return static_fail("ncside * ncside = " + std::to_string((long long)ncside * ncside) +
                           " is outside [1, PARSIM_STATIC_MAX_CELLS = " + std::to_string(kStaticMaxCells) +
                           "]: use a parsim_ctx for larger grids");
    if (n_part < 0 || n_part > kStaticMaxParticles)
        return static_fail("n_part = " + std::to_string(n_part) + " is outside [0, PARSIM_STATIC_MAX_PARTICLES = " +
                           std::to_string(kStaticMaxParticles) + "]: use a parsim_ctx for larger sets");
    if (!d_mx || !d_my || !d_mass) return static_fail("an output pointer is NULL");
    if (n_part > 0 && (!d_x || !d_y || !d_m)) return static_fail("an input pointer is NULL");

    // set-up, once per device with that device current: the array's name is a host-side stand-in
    // in host code, the kernels need its device address
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static void *scratch_of[kMaxDevices] = {};
    int dev = -1;
    cudaError_t rc = cudaGetDevice(&dev);
    if (rc != cudaSuccess) return static_fail(std::string("cudaGetDevice: ") + cudaGetErrorString(rc));
    if (dev < 0 || dev >= kMaxDevices) return static_fail("device index out of range");
    void *ptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!scratch_of[dev]) {
            void *p = nullptr;
            rc = cudaGetSymbolAddress((void **)&p, g_static_scratch);
            if (rc != cudaSuccess || !p)
                return static_fail(std::string("cudaGetSymbolAddress: ") + cudaGetErrorString(rc));
            scratch_of[dev] = p;
        }
        ptr = scratch_of[dev];
    }
    StaticScratch *s = static_cast<StaticScratch *>(ptr);
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    Geom g{};
    g.side = side;
    g.cellw = side / (double)ncside;
    g.inv_cellw = 1.0 / g.cellw;
    g.nc = (int)ncside;
    g.row0 = 0;
    g.rows = (int)ncside;
    g.ncells = g.rows * g.nc;
    g.cell_base = 0;
    const int n = (int)n_part, ncells = g.ncells;
    const int tb = 256;
    const unsigned cell_blocks = (unsigned)((ncells + tb - 1) / tb);
    const unsigned part_blocks = (unsigned)((n + tb - 1) / tb);

    k_static_clear<<<cell_blocks, tb, 0, st>>>(ncells, s->cursor);
    if (n > 0) k_static_count<<<part_blocks, tb, 0, st>>>(g, n, d_x, d_y, s->cursor);
    k_static_scan<<<1, 1024, 0, st>>>(ncells, s->cursor, s->cell_start);
    if (n > 0) {
        k_static_fill<<<part_blocks, tb, 0, st>>>(g, n, d_x, d_y, s->cursor, s->bag);
        k_static_order<<<part_blocks, tb, 0, st>>>(g, n, d_x, d_y, s->cell_start, s->bag, s->sorted);
    }
    // the lane count of the context path (rank_create): from the mean occupancy
    const double ppc = (double)n_part / ((double)ncside * (double)ncside);
    const int lanes = ppc <= 6 ? 4 : ppc <= 12 ? 8 : ppc <= 24 ? 16 : 32;
    const unsigned com_blocks = (unsigned)(((long long)ncells * lanes + kThreads - 1) / kThreads);
#define PS_STATIC_COM(L) \
    k_static_com<L><<<com_blocks, kThreads, 0, st>>>(ncells, d_x, d_y, d_m, s->cell_start, s->sorted, d_mx, d_my, d_mass)
    switch (lanes) {
    case 4: PS_STATIC_COM(4); break;
    case 8: PS_STATIC_COM(8); break;
    case 16: PS_STATIC_COM(16); break;
    default: PS_STATIC_COM(32); break;
    }
#undef PS_STATIC_COM
    rc = cudaGetLastError();
    if (rc != cudaSuccess) return static_fail(std::string("kernel launch: ") + cudaGetErrorString(rc));
    return 0;
}
