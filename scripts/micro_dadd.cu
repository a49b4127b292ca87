// Dependent-issue latency of DADD on one warp: bounds the exact-order list walk when few lists are walked.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/micro_dadd scripts/micro_dadd.cu && /tmp/micro_dadd
#include <cstdio>
#include <cuda_runtime.h>
__global__ void chain(const double* x, double* out, long long* cycles, int n) {
    double v[16];
    for (int i = 0; i < 16; ++i) v[i] = x[i * 32 + threadIdx.x];
    double s = 0.0;
    const long long t0 = clock64();
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) s = __dadd_rn(s, v[i]);
    }
    const long long t1 = clock64();
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cycles = t1 - t0;
}
int main() {
    double *x, *o; long long* c;
    cudaMalloc(&x, 16 * 32 * 8); cudaMalloc(&o, 32 * 8); cudaMalloc(&c, 8);
    cudaMemset(x, 0, 16 * 32 * 8);
    for (int rep = 0; rep < 2; ++rep) {
        chain<<<1, 32>>>(x, o, c, 1024);
        long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("dependent DADD: %.2f cycles each\n", (double)h / (1024.0 * 16));
    }
    return 0;
}
