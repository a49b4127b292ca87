// Timing of the radix sort alone (csrc/k3_onesweep.cu) on one bucket's worth of (key, duration) words.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iapmbackend_b200/csrc -Iinclude [-DOS_MICRO_NO_LOOKBACK]
//      -o scripts/micro_sort scripts/micro_sort.cu && scripts/micro_sort [n] [keys] [vbits]
// OS_MICRO_NO_LOOKBACK skips the wait on predecessor tiles (wrong order, same traffic): the difference is the chain's cost.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#include "../apmbackend_b200/csrc/k3_onesweep.cu"

static uint64_t sm64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

int main(int argc, char** argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 1666667;
    const uint32_t keys = argc > 2 ? atoi(argv[2]) : 10000;
    const int vbits = argc > 3 ? atoi(argv[3]) : 17;
    int kbits = 0;
    while ((1u << kbits) < keys) ++kbits;
    std::vector<uint64_t> h(n);
    uint64_t s = 42;
    for (uint64_t i = 0; i < n; ++i) {
        const double u = (double)(sm64(s) >> 11) / 9007199254740992.0;
        const uint32_t key = (uint32_t)(u * u * u * keys);  // skewed towards the low ids
        const uint32_t val = (uint32_t)(sm64(s) & ((1u << vbits) - 1u));
        h[i] = ((uint64_t)key << 32) | val;
    }
    uint64_t *a, *b, *src;
    unsigned char* scratch;
    const size_t sb = apm_onesweep_scratch_bytes(n);
    cudaMalloc(&a, n * 8); cudaMalloc(&b, n * 8); cudaMalloc(&src, n * 8); cudaMalloc(&scratch, sb);
    cudaMemcpy(src, h.data(), n * 8, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t err;
    uint64_t* res = nullptr;
    float total = 0;
    const int reps = 20;
    for (int r = 0; r < reps + 3; ++r) {
        cudaMemcpyAsync(a, src, n * 8, cudaMemcpyDeviceToDevice, 0);
        cudaEventRecord(e0, 0);
        res = apm_onesweep_sort(0, a, b, n, scratch, sb, vbits, kbits, nullptr, &err);
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r >= 3) total += ms;
    }
    if (!res || err != cudaSuccess) { printf("sort failed: %s\n", cudaGetErrorString(err)); return 1; }
    std::vector<uint64_t> out(n);
    cudaMemcpy(out.data(), res, n * 8, cudaMemcpyDeviceToHost);
    std::sort(h.begin(), h.end());
    const bool ok = out == h;
    printf("n=%llu keys=%u bits=%d+%d: %.1f us per sort (%s)\n", (unsigned long long)n, keys, kbits, vbits, 1000.0 * total / reps,
           ok ? "sorted" : "NOT sorted");
    return 0;
}
