// Counter-based synthetic transaction stream (bench + tests).  Integer-only so the CUDA
// generator and the host generator produce identical columns for the same parameters.
// Shape follows SURVEY.md 8(d): Zipf service popularity (CDF table), per-service duration
// scale times a lognormal quantile table, near-sorted event time with bounded jitter and an
// optional duration shift on some services in the last third of the span (fires signals).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define SYN_HD __host__ __device__ __forceinline__
#else
#define SYN_HD inline
#endif

struct synth_params {
    uint64_t seed;
    uint64_t n;            // transactions in the whole stream
    uint64_t first;        // index of the first transaction to generate (streams can be cut in chunks)
    uint32_t num_services;
    uint32_t num_jvms;
    int64_t t0_ms;
    int64_t span_ms;       // event time of transaction i is t0 + i * span / n + jitter
    uint32_t jitter_ms;    // uniform in [0, jitter_ms]
    uint32_t elapsed_max;  // clamp
    uint32_t shift_mod;    // services with id % shift_mod == 0 run shift_mul x slower in the last third (0 = off)
    uint32_t shift_mul;
    uint32_t q_size;       // entries of the quantile table (power of two)
    uint32_t pad_;
    const uint32_t* cdf;   // [num_services] P(service <= k) * 2^32, non-decreasing, last = 0xFFFFFFFF
    const uint32_t* base;  // [num_services] duration scale (ms)
    const uint32_t* q;     // [q_size] lognormal quantiles * 1024
};

SYN_HD uint64_t syn_mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

SYN_HD void syn_tx(const synth_params& p, uint64_t i, uint32_t* svc, uint32_t* jvm, int64_t* ts, int32_t* el) {
    const uint64_t h1 = syn_mix(p.seed ^ (i * 0xD1B54A32D192ED03ull));
    const uint64_t h2 = syn_mix(h1 ^ 0xA0761D6478BD642Full);
    const uint32_t u = (uint32_t)(h1 >> 32);
    uint32_t lo = 0, hi = p.num_services - 1;  // first k with u <= cdf[k]
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (p.cdf[mid] < u) lo = mid + 1; else hi = mid;
    }
    const uint32_t s = lo;
    *svc = s;
    *jvm = (uint32_t)((h1 & 0xFFFFFFFFull) % p.num_jvms);
    const uint64_t t_off = (uint64_t)(((unsigned __int128)i * (uint64_t)p.span_ms) / p.n);
    const uint32_t jit = p.jitter_ms ? (uint32_t)((h2 >> 40) % (p.jitter_ms + 1)) : 0;
    *ts = p.t0_ms + (int64_t)t_off + jit;
    uint64_t e = ((uint64_t)p.base[s] * p.q[(uint32_t)h2 & (p.q_size - 1)]) >> 10;
    if (p.shift_mod && (s % p.shift_mod) == 0 && 3 * t_off > 2 * (uint64_t)p.span_ms) e *= p.shift_mul;
    if (e < 1) e = 1;
    if (e > p.elapsed_max) e = p.elapsed_max;
    *el = (int32_t)e;
}
