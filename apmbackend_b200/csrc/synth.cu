// libapm_synth.so: synthetic stream generator (device + host twins) and the few raw device
// memory helpers the ctypes tests and bench.py need.  Bench/test support, not product API.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/apm_b200.h"
#include "synth_common.h"

namespace {
__global__ void __launch_bounds__(256) synth_kernel(const synth_params p, uint64_t count, uint32_t* svc, uint32_t* jvm,
                                                    int64_t* ts, int32_t* el) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += stride)
        syn_tx(p, p.first + j, svc + j, jvm + j, ts + j, el + j);
}

// ring pre-fill: a plausible history for every (series, key) column - the level follows the
// service's duration scale (mean ~ 1.13 x base, p95 ~ 2.28 x, p99 ~ 3.2 x for lognormal(0, 0.5)),
// plus a small deterministic noise, all in tenths like the values toFixed(1) produces
__host__ __device__ __forceinline__ double ring_fill_value(uint64_t i, uint64_t ncol, uint64_t seed, const uint32_t* base,
                                                           uint32_t num_services, uint32_t K) {
    const uint64_t col = i % ncol;
    const uint32_t series = (uint32_t)(col / K), key = (uint32_t)(col % K);
    const uint32_t mult = series == 0 ? 113u : (series == 1 ? 228u : 320u);
    const uint64_t hn = syn_mix(seed ^ (i * 0xD1B54A32D192ED03ull) ^ 0x5555);
    const uint32_t tenths = base[key % num_services] * mult / 10u + (uint32_t)(hn % 41u);
    return (double)tenths / 10.0;
}

__global__ void __launch_bounds__(256) ring_fill_kernel(double* out, uint64_t rows, uint64_t ncol, uint64_t seed,
                                                        const uint32_t* base, uint32_t num_services, uint32_t K) {
    const uint64_t total = rows * ncol, stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride)
        out[i] = ring_fill_value(i, ncol, seed, base, num_services, K);
}
}  // namespace

extern "C" {

int synth_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}
int synth_set_device(int d) { return (int)cudaSetDevice(d); }
int synth_malloc(void** p, uint64_t bytes) { return (int)cudaMalloc(p, bytes); }
int synth_free(void* p) { return (int)cudaFree(p); }
int synth_malloc_host(void** p, uint64_t bytes) { return (int)cudaMallocHost(p, bytes); }
int synth_free_host(void* p) { return (int)cudaFreeHost(p); }
int synth_h2d(void* d, const void* h, uint64_t bytes) { return (int)cudaMemcpy(d, h, bytes, cudaMemcpyHostToDevice); }
int synth_d2h(void* h, const void* d, uint64_t bytes) { return (int)cudaMemcpy(h, d, bytes, cudaMemcpyDeviceToHost); }
int synth_sync(void) { return (int)cudaDeviceSynchronize(); }
const char* synth_error_string(int e) { return cudaGetErrorString((cudaError_t)e); }

// tables in p must be device pointers
int synth_generate_device(const synth_params* p, uint64_t count, apm_tx_columns* cols) {
    synth_kernel<<<148 * 8, 256>>>(*p, count, cols->service_id, cols->jvm_id, cols->end_ts_ms, cols->elapsed_ms);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    return (int)cudaDeviceSynchronize();
}

// tables in p must be host pointers
void synth_generate_host(const synth_params* p, uint64_t count, apm_tx_columns* cols) {
    for (uint64_t j = 0; j < count; ++j)
        syn_tx(*p, p->first + j, cols->service_id + j, cols->jvm_id + j, cols->end_ts_ms + j, cols->elapsed_ms + j);
}

// base: device pointer to the stream's duration-scale table
int synth_ring_fill_device(double* out, uint64_t rows, uint64_t ncol, uint64_t seed, const uint32_t* base, uint32_t num_services,
                           uint32_t K) {
    ring_fill_kernel<<<148 * 8, 256>>>(out, rows, ncol, seed, base, num_services, K);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    return (int)cudaDeviceSynchronize();
}

// the same values for a selection of columns, packed [rows][ncols_sel]; base: host pointer
void synth_ring_fill_host(double* out, uint64_t rows, uint64_t ncol, uint64_t seed, const uint32_t* base, uint32_t num_services,
                          uint32_t K, const uint64_t* cols, uint64_t ncols_sel) {
    for (uint64_t r = 0; r < rows; ++r)
        for (uint64_t c = 0; c < ncols_sel; ++c)
            out[r * ncols_sel + c] = ring_fill_value(r * ncol + cols[c], ncol, seed, base, num_services, K);
}

// L2 flush helper for the bench: overwrite a buffer larger than L2
int synth_memset(void* p, int v, uint64_t bytes) { return (int)cudaMemset(p, v, bytes); }

}  // extern "C"
