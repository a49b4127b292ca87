// Design probe (not product): shared/global atomic throughput on B200 for the keyed reduction.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o micro_atomics micro_atomics.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <vector>

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(256) k(const uint32_t* __restrict__ keys, uint64_t n, uint32_t* gtab32, unsigned long long* gtab64, int T) {
    extern __shared__ uint32_t sm[];
    uint32_t* tag = sm;                        // [T]
    uint32_t* c32 = sm + T;                    // [T]
    unsigned long long* c64 = (unsigned long long*)(sm + 2 * T);  // [T]
    unsigned long long* q64 = c64 + T;         // [T]
    for (int i = threadIdx.x; i < T; i += blockDim.x) { tag[i] = 0xFFFFFFFFu; c32[i] = 0; c64[i] = 0; q64[i] = 0; }
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t key = __ldg(keys + i);
        const uint32_t v = (key * 2654435761u) >> 20;
        if (MODE == 0) { acc += key; }                                             // load only
        if (MODE == 1) { atomicAdd(&c32[mix(key) & (T - 1)], 1u); }                 // smem u32 add
        if (MODE == 2) { atomicAdd(&c64[mix(key) & (T - 1)], (unsigned long long)v); }  // smem u64 add
        if (MODE == 3) {                                                            // CAS probe + 3 adds
            uint32_t h = mix(key) & (T - 1);
            for (int p = 0; p < 8; ++p) {
                uint32_t cur = tag[h];
                if (cur != key) { if (cur == 0xFFFFFFFFu) cur = atomicCAS(&tag[h], 0xFFFFFFFFu, key); if (cur != 0xFFFFFFFFu && cur != key) { h = (h + 1) & (T - 1); continue; } }
                atomicAdd(&c32[h], 1u); atomicAdd(&c64[h], (unsigned long long)v); atomicAdd(&q64[h], (unsigned long long)v * v);
                break;
            }
        }
        if (MODE == 4) { atomicAdd(&gtab32[key], 1u); }                             // global RED u32
        if (MODE == 5) { atomicAdd(&gtab32[key], 1u); atomicAdd(&gtab64[key], (unsigned long long)v); atomicAdd(&gtab64[key + 16384], (unsigned long long)v * v); atomicMin(&gtab32[key + 16384], v); atomicMax(&gtab32[key + 32768], v); }
        if (MODE == 6) {                                                            // match_any aggregate then smem u32 add
            uint32_t peers = __match_any_sync(__activemask(), key);
            if ((__ffs(peers) - 1) == (threadIdx.x & 31)) atomicAdd(&c32[mix(key) & (T - 1)], __popc(peers));
        }
        if (MODE == 8) {                                                            // CAS probe + u32-only adds with carries
            uint32_t h = mix(key) & (T - 1);
            uint32_t* lo = (uint32_t*)c64; uint32_t* hi = (uint32_t*)q64;
            for (int p = 0; p < 8; ++p) {
                uint32_t cur = tag[h];
                if (cur != key) { if (cur == 0xFFFFFFFFu) cur = atomicCAS(&tag[h], 0xFFFFFFFFu, key); if (cur != 0xFFFFFFFFu && cur != key) { h = (h + 1) & (T - 1); continue; } }
                atomicAdd(&c32[h], 1u);
                uint32_t o = atomicAdd(&lo[2 * h], v); if (o + v < o) atomicAdd(&lo[2 * h + 1], 1u);
                uint32_t q = v * v; uint32_t o2 = atomicAdd(&hi[2 * h], q); if (o2 + q < o2) atomicAdd(&hi[2 * h + 1], 1u);
                break;
            }
        }
        if (MODE == 7) {                                                            // packed: one u64 add (cnt<<40|sum) + one u64 add sq
            uint32_t h = mix(key) & (T - 1);
            atomicAdd(&c64[h], (1ull << 40) | v); atomicAdd(&q64[h], (unsigned long long)v * v);
        }
    }
    if (MODE == 0 && acc == 0x12345) gtab32[0] = acc;
    __syncthreads();
    if (MODE != 0 && MODE != 4 && MODE != 5) for (int i = threadIdx.x; i < T; i += blockDim.x) if (c32[i] | c64[i]) atomicAdd(&gtab32[i & 1023], c32[i] + (uint32_t)c64[i]);
}

template <int MODE>
float run(const uint32_t* keys, uint64_t n, uint32_t* g32, unsigned long long* g64, int T, int blocks) {
    size_t smem = (size_t)T * (4 + 4 + 8 + 8);
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<MODE><<<blocks, 256, smem>>>(keys, n, g32, g64, T);
    cudaEventRecord(a);
    for (int r = 0; r < 3; ++r) k<MODE><<<blocks, 256, smem>>>(keys, n, g32, g64, T);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    cudaError_t e = cudaGetLastError(); if (e) printf("err %s\n", cudaGetErrorString(e));
    return ms / 3;
}

int main() {
    const uint64_t n = 1ull << 26;
    const int K = 10000;
    std::vector<uint32_t> hz(n), hu(n);
    // Zipf(1.0) via inverse CDF table
    std::vector<double> cdf(K); double s = 0; for (int i = 0; i < K; ++i) { s += 1.0 / (i + 1); cdf[i] = s; }
    uint64_t st = 88172645463325252ull;
    for (uint64_t i = 0; i < n; ++i) {
        st ^= st << 13; st ^= st >> 7; st ^= st << 17;
        double u = (st >> 11) * (1.0 / 9007199254740992.0) * s;
        int lo = 0, hi = K - 1; while (lo < hi) { int m = (lo + hi) / 2; if (cdf[m] < u) lo = m + 1; else hi = m; }
        hz[i] = lo; hu[i] = (uint32_t)(st >> 33) % K;
    }
    uint32_t *dz, *du, *g32; unsigned long long* g64;
    cudaMalloc(&dz, n * 4); cudaMalloc(&du, n * 4); cudaMalloc(&g32, 1 << 20); cudaMalloc(&g64, 1 << 20);
    cudaMemcpy(dz, hz.data(), n * 4, cudaMemcpyHostToDevice); cudaMemcpy(du, hu.data(), n * 4, cudaMemcpyHostToDevice);
    cudaMemset(g32, 0, 1 << 20); cudaMemset(g64, 0, 1 << 20);
    const char* names[] = {"load only", "smem u32 add", "smem u64 add", "smem CAS-probe + u32 + 2x u64 add", "global RED u32", "global 5 atomics (v1 kernel)", "match_any + smem u32", "smem packed 2x u64 add", "smem CAS-probe + 3x u32 add w/ carry"};
    for (int T : {2048, 4096}) for (int occ : {2, 4}) {
        int blocks = 148 * occ;
        printf("--- T=%d blocks=%d\n", T, blocks);
        for (int dist = 0; dist < 2; ++dist) {
            const uint32_t* keys = dist ? dz : du;
            float ms[9];
            ms[0] = run<0>(keys, n, g32, g64, T, blocks); ms[1] = run<1>(keys, n, g32, g64, T, blocks); ms[2] = run<2>(keys, n, g32, g64, T, blocks);
            ms[3] = run<3>(keys, n, g32, g64, T, blocks); ms[4] = run<4>(keys, n, g32, g64, T, blocks); ms[5] = run<5>(keys, n, g32, g64, T, blocks);
            ms[6] = run<6>(keys, n, g32, g64, T, blocks); ms[7] = run<7>(keys, n, g32, g64, T, blocks); ms[8] = run<8>(keys, n, g32, g64, T, blocks);
            for (int m = 0; m < 9; ++m) printf("%-8s %-36s %8.3f ms  %7.2f Gtx/s\n", dist ? "zipf" : "uniform", names[m], ms[m], n / ms[m] / 1e6);
        }
    }
    return 0;
}
