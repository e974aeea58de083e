// Micro-benchmark: does the NUMBER of concurrent DRAM streams limit a copy-like kernel on B200?
// soa5: 5 input arrays -> 5 output arrays (the layout k_move works on); aos: one 48-byte record
// array -> one record array; plain: one double2 array -> one double2 array.
#include <cstdio>
#include <cuda_runtime.h>
struct __align__(16) Rec { double2 p, v; double m; unsigned id; int cell; };
__global__ void soa5(int n, const double2 *p, const double2 *v, const double *m, const unsigned *id, const int *c,
                     double2 *po, double2 *vo, double *mo, unsigned *ido, int *co) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { double2 a = p[i], b = v[i]; double mm = m[i]; unsigned d = id[i]; int e = c[i];
        po[i] = a; vo[i] = b; mo[i] = mm; ido[i] = d; co[i] = e; }
}
__global__ void aos(int n, const Rec *in, Rec *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const double2 *q = reinterpret_cast<const double2 *>(in + i);
        double2 a = q[0], b = q[1], c = q[2];
        double2 *o = reinterpret_cast<double2 *>(out + i);
        o[0] = a; o[1] = b; o[2] = c;
    }
}
__global__ void plain(long long n, const double2 *in, double2 *out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i];
}
int main() {
    const int n = 100000000;
    double2 *p, *v, *po, *vo; double *m, *mo; unsigned *id, *ido; int *c, *co; Rec *ri, *ro;
    cudaMalloc(&p, 16ll * n); cudaMalloc(&v, 16ll * n); cudaMalloc(&po, 16ll * n); cudaMalloc(&vo, 16ll * n);
    cudaMalloc(&m, 8ll * n); cudaMalloc(&mo, 8ll * n); cudaMalloc(&id, 4ll * n); cudaMalloc(&ido, 4ll * n);
    cudaMalloc(&c, 4ll * n); cudaMalloc(&co, 4ll * n); cudaMalloc(&ri, 48ll * n); cudaMalloc(&ro, 48ll * n);
    cudaMemset(p, 0, 16ll * n); cudaMemset(v, 0, 16ll * n); cudaMemset(m, 0, 8ll * n); cudaMemset(id, 0, 4ll * n);
    cudaMemset(c, 0, 4ll * n); cudaMemset(ri, 0, 48ll * n);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float ms;
    const int g = (n + 255) / 256;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(a); for (int k = 0; k < 5; k++) soa5<<<g, 256>>>(n, p, v, m, id, c, po, vo, mo, ido, co);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("soa5 : %.3f ms  %.0f GB/s\n", ms / 5, 96.0 * n / (ms / 5 * 1e-3) / 1e9);
        cudaEventRecord(a); for (int k = 0; k < 5; k++) aos<<<g, 256>>>(n, ri, ro);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("aos  : %.3f ms  %.0f GB/s\n", ms / 5, 96.0 * n / (ms / 5 * 1e-3) / 1e9);
        const long long n2 = 3ll * n; const int g2 = (int)((n2 + 255) / 256);
        cudaEventRecord(a); for (int k = 0; k < 5; k++) plain<<<g2, 256>>>(n2, (const double2 *)ri, (double2 *)ro);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("plain: %.3f ms  %.0f GB/s\n", ms / 5, 96.0 * n / (ms / 5 * 1e-3) / 1e9);
    }
    return 0;
}
