// Shared device/host definitions of libapm_b200 (sm_100a only).
#pragma once
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/apm_b200.h"

#define APM_MAX_SLOTS 64  // window + buffer + 1 must not exceed this
#define APM_MAX_DEVICES 64

// SMs of the current device (148 on B200), asked once per device ordinal: grids are sized in multiples of it.
int apm_sm_count();
// True exactly once per device ordinal and `done` mask: kernel attributes such as the dynamic shared-memory
// limit belong to the device that is current when they are set, so a process that holds contexts on several
// GPUs has to set them on each.
bool apm_first_on_device(unsigned long long* done);

// Loads one kernel on the current device (see apm_preload_kernels).  APM_SMEM_CARVEOUT=<percent> additionally gives
// every kernel of the library the same preferred shared-memory carve-out, for A/B runs of small shapes where kernels
// with very different shared-memory needs alternate on the same SMs.
inline cudaError_t apm_preload_one(cudaFuncAttributes* a, const void* f) {
    const cudaError_t e = cudaFuncGetAttributes(a, f);
    if (e != cudaSuccess) return e;
    static const int carveout = [] { const char* v = getenv("APM_SMEM_CARVEOUT"); return v ? atoi(v) : -1; }();
    if (carveout >= 0) return cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
    return cudaSuccess;
}

// Makes `device` current for the duration of an entry point and restores the caller's device afterwards:
// one host process may drive a context per GPU (the reference stage is one process, apm_manager.js:329-347).
struct ApmDeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit ApmDeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); }
        if (prev != device) switched = cudaSetDevice(device) == cudaSuccess;
    }
    ~ApmDeviceGuard() { if (switched && prev >= 0) cudaSetDevice(prev); }
    ApmDeviceGuard(const ApmDeviceGuard&) = delete;
    ApmDeviceGuard& operator=(const ApmDeviceGuard&) = delete;
};
#define APM_ENTER(ctx) ApmDeviceGuard apm_device_guard_((ctx)->cfg.device)

// (key, bucket-slot) accumulators: everything addData (stream_calc_stats.js:115-134) would
// have appended to buckets[label], reduced to order-independent integers.  Structure of
// arrays, each [slot][key], so the per-key tick kernel reads them coalesced and every update
// is a fire-and-forget L2 reduction:
//   sum   = sum of elapsed                       (u64: < 2^32 values of < 2^31)
//   sq_lo = sum of (elapsed^2 & 0xFFFFFFFF)      (u64: cannot overflow below 2^32 updates)
//   sq_hi = sum of (elapsed^2 >> 32)             (u64; only durations >= 2^16 ms touch it)
// so the exact sum of squares is sq_hi * 2^32 + sq_lo without any carry between atomics.
struct AccSoA {
    unsigned long long* sum;
    unsigned long long* sq_lo;
    unsigned long long* sq_hi;
    uint32_t* cnt;
    uint32_t* vmin;  // 0xFFFFFFFF when empty
    uint32_t* vmax;
    uint32_t* run_start;  // first word of the key's run in the slot's sorted log (percentile mode)
};
#define APM_ACC_BYTES_PER_ENTRY 40

// Per bucket-slot bookkeeping of the value log (device resident).
struct SlotMeta {
    unsigned long long count;   // entries appended
    unsigned long long sorted;  // entries covered by the last sort
    uint32_t max_value;         // largest elapsed appended (bounds the radix passes)
    uint32_t pad_;
};

// Device counters, one cache line.
struct DevCounters {
    unsigned long long ingested, dropped, invalid, future, log_overflow;
    unsigned long long inc_fallbacks;  // guarded incremental baseline: columns that had to be re-summed exactly
    uint32_t max_value;      // largest elapsed ever ingested (bounds the radix passes)
    uint32_t unseen_records; // record slots of keys not seen yet, over the retained ticks (0: every retained record is valid)
    uint32_t pad_;
    uint32_t pending_lists;  // keys that hold a loaded lag list but have not emitted yet (may still need realigning)
};

struct WindowSlots {
    int n;
    int slot[APM_MAX_SLOTS];
};

#define CUDA_TRY(ctx, expr)                                                                       \
    do {                                                                                          \
        cudaError_t e_ = (expr);                                                                  \
        if (e_ != cudaSuccess) return apm_fail_cuda((ctx), e_, #expr, __FILE__, __LINE__);        \
    } while (0)

struct apm_ctx;
apm_status apm_fail_cuda(apm_ctx* ctx, cudaError_t e, const char* what, const char* file, int line);
apm_status apm_fail(apm_ctx* ctx, apm_status st, const std::string& msg);

// ---------------------------------------------------------------------------------------------
// Exact helpers shared by host and device
// ---------------------------------------------------------------------------------------------

// generateBucketLabel (stream_calc_stats.js:89-91): floor(ts / bucket_ms) for ts >= 0.
__host__ __device__ __forceinline__ int64_t apm_label(int64_t ts, int64_t bucket_ms, double inv_bucket) {
#ifdef __CUDA_ARCH__
    // ts < 2^53: one multiply and a one-step fix-up is exact and avoids the 64-bit divide sequence
    int64_t q = (int64_t)(__dmul_rn((double)ts, inv_bucket));
    int64_t r = ts - q * bucket_ms;
    if (r < 0) { q -= 1; r += bucket_ms; }
    if (r >= bucket_ms) { q += 1; }
    return q;
#else
    (void)inv_bucket;
    return ts / bucket_ms;
#endif
}

// parseFloat(x.toFixed(digits)) for x >= 0 or NaN, digits in {1, 2}: what the next stage reads
// after the CSV hop (entries.js:65-73 -> 58-61).  toFixed picks the integer n closest to
// x * 10^digits on the EXACT binary value of x, the larger n on a tie; strtod of the decimal
// string is the correctly rounded n / 10^digits, which one IEEE division reproduces
// (n < 2^53).  Negative inputs (never produced by the stats stage) take the mirrored path.
__host__ __device__ __forceinline__ double apm_quantize_fixed(double x, int digits) {
    if (x != x) return x;
    const bool neg = x < 0;
    if (neg) x = -x;
    if (x >= 4.0e15) return neg ? -x : x;  // already an integer multiple far above 0.01
    unsigned long long bits;
#ifdef __CUDA_ARCH__
    bits = (unsigned long long)__double_as_longlong(x);
#else
    __builtin_memcpy(&bits, &x, 8);
#endif
    const int be = (int)((bits >> 52) & 0x7FF);
    unsigned long long m = bits & 0xFFFFFFFFFFFFFull;
    int e;
    if (be == 0) { e = -1074; } else { m |= 1ull << 52; e = be - 1075; }  // x = m * 2^e
    const unsigned long long p = digits == 1 ? 10ull : 100ull;
    unsigned long long v = m * p;  // < 2^60
    unsigned long long n;
    if (e >= 0) {
        n = v << e;  // x < 4e15 keeps this below 2^63
    } else if (-e >= 64) {
        n = 0;
    } else {
        const int s = -e;
        const unsigned long long q = v >> s;
        const unsigned long long r = v - (q << s);
        n = q + ((r >= (1ull << (s - 1))) ? 1ull : 0ull);
    }
    const double d = digits == 1 ? 10.0 : 100.0;
#ifdef __CUDA_ARCH__
    const double y = __ddiv_rn((double)n, d);
#else
    const double y = (double)n / d;
#endif
    return neg ? -y : y;
}
