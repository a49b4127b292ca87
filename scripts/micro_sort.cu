// Timing and checking of the bucket sorts alone on one bucket's worth of (key, duration) words:
//   onesweep (csrc/k3_onesweep.cu): LSD radix, one look-back pass per 8-bit digit
//   group    (csrc/k3_groupsort.cu): partition by the previous bucket's splitters + shared-memory group sorts
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -Iapmbackend_b200/csrc -Iinclude [-DOS_MICRO_NO_LOOKBACK]
//      -o scripts/micro_sort scripts/micro_sort.cu && scripts/micro_sort [n] [keys] [mode]
// mode: zipf (Zipf(1) key popularity, per-key lognormal-ish durations: the bench's shape), cube (u^3 keys, uniform values),
//       shift (the plan comes from a bucket whose durations were 5x shorter), equal (one key, one value), tiny.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>
#include <mutex>
#include <cuda_runtime.h>
#include "../apmbackend_b200/csrc/k3_onesweep.cu"
#include "../apmbackend_b200/csrc/k3_groupsort.cu"

int apm_sm_count() { int n = 148; cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, 0); return n; }
bool apm_first_on_device(unsigned long long* done) { if (*done) return false; *done = 1; return true; }
apm_status apm_fail_cuda(apm_ctx*, cudaError_t, const char*, const char*, int) { return APM_E_DEVICE; }
apm_status apm_fail(apm_ctx*, apm_status st, const std::string&) { return st; }

static uint64_t sm64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static std::vector<uint64_t> make_bucket(uint64_t n, uint32_t keys, const char* mode, uint64_t seed, double scale) {
    std::vector<uint64_t> h(n);
    uint64_t s = seed;
    std::vector<double> cdf(keys);
    double acc = 0;
    for (uint32_t k = 0; k < keys; ++k) { acc += 1.0 / (k + 1); cdf[k] = acc; }
    for (uint64_t i = 0; i < n; ++i) {
        const double u = (double)(sm64(s) >> 11) / 9007199254740992.0;
        uint32_t key, val;
        if (!strcmp(mode, "cube")) {
            key = (uint32_t)(u * u * u * keys);
            val = (uint32_t)(sm64(s) & 0x1FFFFu);
        } else if (!strcmp(mode, "equal")) {
            key = 3; val = 777;
        } else {
            key = (uint32_t)(std::lower_bound(cdf.begin(), cdf.end(), u * acc) - cdf.begin());
            if (key >= keys) key = keys - 1;
            const double g = (double)(sm64(s) >> 11) / 9007199254740992.0, g2 = (double)(sm64(s) >> 11) / 9007199254740992.0;
            const double z = sqrt(-2.0 * log(g + 1e-300)) * cos(6.283185307179586 * g2);
            const double base = 20 + (key * 2654435761u) % 400;
            double v = base * exp(0.5 * z) * scale;
            val = (uint32_t)(v < 1 ? 1 : (v > 60000 ? 60000 : v));
        }
        h[i] = ((uint64_t)key << 32) | val;
    }
    return h;
}

int main(int argc, char** argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 1666667;
    const uint32_t keys = argc > 2 ? atoi(argv[2]) : 10000;
    const char* mode = argc > 3 ? argv[3] : "zipf";
    int kbits = 0;
    while ((1u << kbits) < keys) ++kbits;
    const bool shift = !strcmp(mode, "shift");
    std::vector<uint64_t> prev = make_bucket(n, keys, shift ? "zipf" : mode, 41, shift ? 0.2 : 1.0);
    std::vector<uint64_t> h = make_bucket(n, keys, shift ? "zipf" : mode, 42, 1.0);
    uint32_t maxv = 1;
    for (uint64_t x : h) maxv = std::max(maxv, (uint32_t)x);
    for (uint64_t x : prev) maxv = std::max(maxv, (uint32_t)x);
    int vbits = 0;
    while ((1ull << vbits) <= maxv) ++vbits;
    uint64_t *a, *b, *src, *psorted;
    unsigned char *scratch, *gsc;
    const size_t sb = apm_onesweep_scratch_bytes(n);
    cudaMalloc(&a, n * 8); cudaMalloc(&b, n * 8); cudaMalloc(&src, n * 8); cudaMalloc(&psorted, n * 8); cudaMalloc(&scratch, sb);
    cudaMalloc(&gsc, apm_groupsort_scratch_bytes(keys));
    cudaMemcpy(src, h.data(), n * 8, cudaMemcpyHostToDevice);
    std::sort(prev.begin(), prev.end());
    cudaMemcpy(psorted, prev.data(), n * 8, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t err = cudaSuccess;
    std::vector<uint64_t> want = h, out(n);
    std::sort(want.begin(), want.end());
    const int reps = 20;
    // Every sort is queued behind a fresh copy of the unsorted bucket and nothing waits in between, as in the engine,
    // where the host runs ahead of the device: launch latency hides behind the previous kernel.  The copies' own time
    // (measured alone) is subtracted.
    float copy_ms = 0;
    {
        for (int r = 0; r < 3; ++r) cudaMemcpyAsync(a, src, n * 8, cudaMemcpyDeviceToDevice, 0);
        cudaEventRecord(e0, 0);
        for (int r = 0; r < reps; ++r) cudaMemcpyAsync(a, src, n * 8, cudaMemcpyDeviceToDevice, 0);
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&copy_ms, e0, e1);
    }
    // ---- onesweep
    {
        uint64_t* res = nullptr;
        float total = 0;
        for (int r = 0; r < reps + 3; ++r) {
            if (r == 3) cudaEventRecord(e0, 0);
            cudaMemcpyAsync(a, src, n * 8, cudaMemcpyDeviceToDevice, 0);
            res = apm_onesweep_sort(0, a, b, n, scratch, sb, vbits, kbits, nullptr, &err);
        }
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&total, e0, e1);
        bool ok = res && err == cudaSuccess;
        if (ok) { cudaMemcpy(out.data(), res, n * 8, cudaMemcpyDeviceToHost); ok = out == want; }
        printf("onesweep n=%llu keys=%u bits=%d+%d mode=%s: %.1f us per sort (%s)\n", (unsigned long long)n, keys, kbits, vbits, mode,
               1000.0 * (total - copy_ms) / reps, ok ? "sorted" : "NOT sorted");
    }
    // ---- group sort: the first bucket is planned from the previous (sorted) bucket, every later one from the plan the
    //      sort before it left behind - the engine's steady state
    {
        apm_groupsort_state gs;
        float total = 0;
        bool took = true;
        err = apm_groupsort_plan_from(0, &gs, gsc, keys, psorted, n, nullptr);
        for (int r = 0; r < reps + 3; ++r) {
            if (r == 3) cudaEventRecord(e0, 0);
            cudaMemcpyAsync(a, src, n * 8, cudaMemcpyDeviceToDevice, 0);
            took = took && apm_groupsort(0, &gs, gsc, keys, a, b, n, vbits, nullptr, &err);
        }
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&total, e0, e1);
        bool ok = took && err == cudaSuccess;
        if (ok) { cudaMemcpy(out.data(), b, n * 8, cudaMemcpyDeviceToHost); ok = out == want; }
        printf("group    n=%llu keys=%u mode=%s: %.1f us per sort incl. the next plan (%s%s), splitters now %u\n",
               (unsigned long long)n, keys, mode, 1000.0 * (total - copy_ms) / reps, took ? "" : "no plan: not taken, ", ok ? "sorted" : "NOT sorted", gs.P);
    }
    printf("last CUDA error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
