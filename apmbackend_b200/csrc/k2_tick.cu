// K2 + K3b: per-key window combine, percentile selection and the toFixed hop.
//
// Replaces removeOldBuckets (stream_calc_stats.js:103-113) with a slot clear and
// generateAllStatsToQueue (:157-203) + calcPercentile (util_methods.js:112-142) +
// StatEntry.toCSVString / re-parse (entries.js:52-73, 179-180) with one thread per key.
#include "ctx.h"

namespace {

__global__ void __launch_bounds__(256) clear_slots_kernel(const AccSoA acc, SlotMeta* __restrict__ meta, uint32_t K,
                                                          unsigned long long slot_mask) {
    const int slot = blockIdx.y;  // grid.y = slot
    if (!((slot_mask >> slot) & 1ull)) return;
    const size_t base = (size_t)slot * K;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)K; i += (size_t)gridDim.x * blockDim.x) {
        acc.sum[base + i] = 0; acc.sq_lo[base + i] = 0; acc.sq_hi[base + i] = 0;
        acc.cnt[base + i] = 0; acc.vmin[base + i] = 0xFFFFFFFFu; acc.vmax[base + i] = 0; acc.run_start[base + i] = 0;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        meta[slot].count = 0;
        meta[slot].sorted = 0;
        meta[slot].max_value = 0;
    }
}

struct Run {
    const uint64_t* p;  // first word of this key's ascending run inside a sorted bucket log
    uint32_t n;
};

// number of values <= v in an ascending run (values are the low 32 bits)
__device__ __forceinline__ uint32_t count_le(const Run& r, uint32_t v) {
    uint32_t lo = 0, hi = r.n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if ((uint32_t)__ldg(r.p + mid) <= v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// the same for a run that was staged in shared memory as plain 32-bit values
struct RunS {
    const uint32_t* p;
    uint32_t n;
};
__device__ __forceinline__ uint32_t count_le(const RunS& r, uint32_t v) {
    uint32_t lo = 0, hi = r.n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (r.p[mid] <= v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// k-th smallest (0-based) of the union of the runs; also returns how many values are <= it
__device__ uint32_t select_kth(const Run* runs, int nruns, uint64_t k, uint32_t vmin, uint32_t vmax, uint64_t* n_le) {
    uint32_t lo = vmin, hi = vmax;  // invariant: answer in [lo, hi]
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        uint64_t c = 0;
        for (int r = 0; r < nruns; ++r) c += count_le(runs[r], mid);
        if (c >= k + 1) hi = mid; else lo = mid + 1;
    }
    uint64_t c = 0;
    for (int r = 0; r < nruns; ++r) c += count_le(runs[r], lo);
    *n_le = c;
    return lo;
}

// smallest value strictly greater than v over all runs (caller guarantees one exists)
__device__ uint32_t next_above(const Run* runs, int nruns, uint32_t v) {
    uint32_t best = 0xFFFFFFFFu;
    for (int r = 0; r < nruns; ++r) {
        const uint32_t i = count_le(runs[r], v);
        if (i < runs[r].n) {
            const uint32_t x = (uint32_t)__ldg(runs[r].p + i);
            best = x < best ? x : best;
        }
    }
    return best;
}

struct TickParams {
    AccSoA acc;
    const uint8_t* seen;
    uint64_t* const* logs;
    const SlotMeta* slot_meta;
    uint32_t K;
    WindowSlots win;
    int npct;
    double pfrac[APM_MAX_PCT];
    int zpct[2];
    double tpm_div;
    int64_t tick_ts;
    apm_stat_record* stats;  // nullptr when stat records are not retained
    double* zin;             // [4][K]
    uint8_t* emit;           // [K]
    uint32_t* unseen_records;
    int with_pct;
    uint64_t log_cap;
};

// calcPercentile's index rule (util_methods.js:112-142), double index arithmetic without contraction:
// the percentile is element i0 of the sorted window, or the mean of elements i0 and i0 + 1 when `pair`
__device__ __forceinline__ void pct_index(double pfrac, uint64_t cnt, uint64_t* i0_out, bool* pair_out) {
    const double n = (double)cnt;
    const double index = __dsub_rn(__dmul_rn(pfrac, n), 1.0);
    uint64_t i0;
    bool pair = false;
    if (pfrac == 0.0) {
        i0 = 0;
    } else if (pfrac == 1.0) {
        i0 = cnt - 1;
    } else if (cnt == 1 || index == trunc(index)) {
        i0 = index < 0 ? 0 : (uint64_t)index;
    } else {
        const double ci = ceil(index);
        i0 = ci < 0 ? 0 : (uint64_t)ci;
        pair = i0 != cnt - 1;
    }
    *i0_out = i0;
    *pair_out = pair;
}

// mean and (extension) population stddev of a window from its exact integer moments
__device__ __forceinline__ void key_moments(uint64_t cnt, uint64_t sum, uint64_t sq_lo, uint64_t sq_hi, double* avg, double* sd) {
    *avg = __ddiv_rn((double)sum, (double)cnt);  // :181, one correctly rounded divide of exact integers
    // var = (cnt*sumsq - sum^2) / cnt^2 in 128-bit integers (domain: result < 2^128)
    const uint64_t a_lo = cnt * sq_lo, a_hi = __umul64hi(cnt, sq_lo) + cnt * sq_hi;
    const uint64_t b_lo = sum * sum, b_hi = __umul64hi(sum, sum);
    const uint64_t d_lo = a_lo - b_lo, d_hi = a_hi - b_hi - (a_lo < b_lo ? 1u : 0u);
    const double num = __dadd_rn(__dmul_rn((double)d_hi, 18446744073709551616.0), (double)d_lo);
    *sd = __ddiv_rn(__dsqrt_rn(num), (double)cnt);
}

// the stat record and what the z-score stage parses out of the `st` line
__device__ __forceinline__ void key_epilogue(const TickParams& p, uint32_t key, uint64_t cnt, uint64_t sum, uint32_t vmin, uint32_t vmax,
                                             double avg, double sd, const double* pct) {
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const double tpm = __ddiv_rn((double)cnt, p.tpm_div);  // :186
    if (p.stats) {
        apm_stat_record r;
        r.tick_ts = p.tick_ts;
        r.key = key;
        r.flags = 1u | (cnt ? 2u : 0u);
        r.count = cnt;
        r.sum = sum;
        r.tpm = tpm;
        r.avg = avg;
#pragma unroll
        for (int q = 0; q < APM_MAX_PCT; ++q) r.pct[q] = pct[q];
        r.min = cnt ? (int32_t)vmin : 0;
        r.max = (int32_t)vmax;
        r.stddev = sd;
        p.stats[key] = r;
    }
    // ---- what the z-score stage parses out of the `st` line (entries.js:71-73 -> 58-61)
    p.zin[key] = apm_quantize_fixed(tpm, 2);
    p.zin[(size_t)p.K + key] = apm_quantize_fixed(avg, 1);
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        const int q = p.zpct[s];
        p.zin[(size_t)(2 + s) * p.K + key] = (q >= 0 && q < p.npct) ? apm_quantize_fixed(pct[q], 1) : nan;
    }
}

// 128-bit accumulate of one slot's sum of squares (stored as sq_hi * 2^32 + sq_lo)
__device__ __forceinline__ void add_sq(uint64_t& sq_lo, uint64_t& sq_hi, uint64_t lo_part, uint64_t hi_part) {
    const uint64_t add_lo = lo_part + (hi_part << 32);
    const uint64_t add_hi = (hi_part >> 32) + (add_lo < lo_part ? 1u : 0u);
    const uint64_t t = sq_lo + add_lo;
    sq_hi += add_hi + (t < sq_lo ? 1u : 0u);
    sq_lo = t;
}

__global__ void __launch_bounds__(128) tick_stats_kernel(const TickParams p) {
    const uint32_t key = blockIdx.x * blockDim.x + threadIdx.x;
    if (key >= p.K) return;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    if (!p.seen[key]) {
        if (p.stats) p.stats[key].flags = 0;
        p.emit[key] = 0;
        atomicAdd(p.unseen_records, 1u);
        return;
    }
    p.emit[key] = 1;
    // ---- window combine (:171-178): exact integer moments
    uint64_t cnt = 0, sum = 0, sq_lo = 0, sq_hi = 0;
    uint32_t vmin = 0xFFFFFFFFu, vmax = 0;
    Run runs[APM_MAX_SLOTS];
    int nruns = 0;
    for (int w = 0; w < p.win.n; ++w) {
        const int slot = p.win.slot[w];
        const size_t ai = (size_t)slot * p.K + key;
        const uint32_t c = p.acc.cnt[ai];
        if (c == 0) continue;
        cnt += c;
        const uint32_t mn = p.acc.vmin[ai], mx = p.acc.vmax[ai];
        vmin = mn < vmin ? mn : vmin;
        vmax = mx > vmax ? mx : vmax;
        sum += p.acc.sum[ai];
        // sum of squares = sq_hi * 2^32 + sq_lo, accumulated as a 128-bit (hi:lo) pair
        add_sq(sq_lo, sq_hi, p.acc.sq_lo[ai], p.acc.sq_hi[ai]);
        if (p.with_pct) {
            // the rebuild of the planes from the sorted log recorded where this key's run starts
            runs[nruns].p = p.logs[slot] + p.acc.run_start[ai];
            runs[nruns].n = c;
            ++nruns;
        }
    }
    double avg = nan, sd = nan, pct[APM_MAX_PCT] = {nan, nan, nan, nan};
    if (cnt != 0) {
        key_moments(cnt, sum, sq_lo, sq_hi, &avg, &sd);
        for (int q = 0; q < (p.with_pct ? p.npct : 0); ++q) {
            uint64_t i0;
            bool pair;
            pct_index(p.pfrac[q], cnt, &i0, &pair);
            uint32_t v0, v1;
            if (nruns == 1) {
                v0 = (uint32_t)__ldg(runs[0].p + i0);
                v1 = pair ? (uint32_t)__ldg(runs[0].p + i0 + 1) : v0;
            } else {
                uint64_t n_le;
                v0 = select_kth(runs, nruns, i0, vmin, vmax, &n_le);
                v1 = v0;
                if (pair && n_le < i0 + 2) v1 = next_above(runs, nruns, v0);
            }
            pct[q] = pair ? __dmul_rn(__dadd_rn((double)v0, (double)v1), 0.5) : (double)v0;
        }
    }
    key_epilogue(p, key, cnt, sum, vmin, vmax, avg, sd, pct);
}

// ---- sliding windows (more than one bucket per window): one warp per key, one lane per bucket.
// The thread-per-key kernel above walks the window's runs one after another inside every step of
// the value bisection - 31 dependent binary searches per step with the reference's settings, and
// with a few hundred keys only a handful of SMs are busy at all.  Here lane w owns the runs of
// window buckets w and w + 32: a bisection step is one binary search per lane and a warp sum.
__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

constexpr int TSW_STAGE = 2560;  // values of one key's window a warp stages in shared memory (4 warps: 40 KB)

__global__ void __launch_bounds__(128) tick_stats_warp_kernel(const TickParams p) {
    const uint32_t key = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (key >= p.K) return;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    if (!p.seen[key]) {
        if (lane == 0) {
            if (p.stats) p.stats[key].flags = 0;
            p.emit[key] = 0;
            atomicAdd(p.unseen_records, 1u);
        }
        return;
    }
    uint64_t cnt = 0, sum = 0, sq_lo = 0, sq_hi = 0;
    uint32_t vmin = 0xFFFFFFFFu, vmax = 0;
    Run runs[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int w = lane + 32 * h;
        runs[h].p = nullptr;
        runs[h].n = 0;
        if (w < p.win.n) {
            const int slot = p.win.slot[w];
            const size_t ai = (size_t)slot * p.K + key;
            const uint32_t c = p.acc.cnt[ai];
            if (c != 0) {
                cnt += c;
                const uint32_t mn = p.acc.vmin[ai], mx = p.acc.vmax[ai];
                vmin = mn < vmin ? mn : vmin;
                vmax = mx > vmax ? mx : vmax;
                sum += p.acc.sum[ai];
                add_sq(sq_lo, sq_hi, p.acc.sq_lo[ai], p.acc.sq_hi[ai]);
                runs[h].p = p.logs[slot] + p.acc.run_start[ai];
                runs[h].n = c;
            }
        }
    }
    // exact integer moments of the whole window: every lane ends with the same totals
    cnt = warp_sum_u64(cnt);
    sum = warp_sum_u64(sum);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint64_t olo = __shfl_xor_sync(0xffffffffu, sq_lo, o), ohi = __shfl_xor_sync(0xffffffffu, sq_hi, o);
        const uint64_t t = sq_lo + olo;
        sq_hi += ohi + (t < sq_lo ? 1u : 0u);
        sq_lo = t;
    }
    vmin = __reduce_min_sync(0xffffffffu, vmin);
    vmax = __reduce_max_sync(0xffffffffu, vmax);
    const int nruns = __popc(__ballot_sync(0xffffffffu, runs[0].n != 0)) + __popc(__ballot_sync(0xffffffffu, runs[1].n != 0));

    double avg = nan, sd = nan, pct[APM_MAX_PCT] = {nan, nan, nan, nan};
    // A window of up to TSW_STAGE values is copied into the warp's shared-memory stretch first: the value bisection below
    // is a chain of ~17 x 2 x 5 dependent loads per percentile, and out of L2 that chain was the whole kernel (22 us for
    // the reference's 400 keys x 31 buckets).
    __shared__ uint32_t s_stage[4][TSW_STAGE];
    const bool staged = p.with_pct && nruns > 1 && cnt <= (uint64_t)TSW_STAGE;
    RunS sruns[2] = {{nullptr, 0}, {nullptr, 0}};
    if (staged) {
        uint32_t* mine = s_stage[(threadIdx.x >> 5) & 3];
        const uint32_t len = runs[0].n + runs[1].n;
        uint32_t incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        uint32_t at = incl - len;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            sruns[h].p = mine + at;
            sruns[h].n = runs[h].n;
            // eight loads in flight per round: one at a time, a hot key's 60-word run cost 60 round trips to L2
            for (uint32_t i = 0; i < runs[h].n; i += 8) {
                uint32_t v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = i + u < runs[h].n ? (uint32_t)__ldg(runs[h].p + i + u) : 0u;
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (i + u < runs[h].n) mine[at + i + u] = v[u];
            }
            at += runs[h].n;
        }
        __syncwarp();
    }
    if (cnt != 0) {
        key_moments(cnt, sum, sq_lo, sq_hi, &avg, &sd);
        for (int q = 0; q < (p.with_pct ? p.npct : 0); ++q) {
            uint64_t i0;
            bool pair;
            pct_index(p.pfrac[q], cnt, &i0, &pair);
            if (nruns == 1) {  // a single populated bucket: the window is that run, index it directly
                uint32_t a = 0, b = 0;
                if (runs[0].n | runs[1].n) {
                    const Run& r = runs[0].n ? runs[0] : runs[1];
                    a = (uint32_t)__ldg(r.p + i0);
                    b = pair ? (uint32_t)__ldg(r.p + i0 + 1) : a;
                }
                const uint32_t v0 = __reduce_max_sync(0xffffffffu, a), v1 = __reduce_max_sync(0xffffffffu, b);
                pct[q] = pair ? __dmul_rn(__dadd_rn((double)v0, (double)v1), 0.5) : (double)v0;
                continue;
            }
            // value bisection: the i0-th smallest is the least v with #(values <= v) >= i0 + 1
            uint32_t lo = vmin, hi = vmax;
            while (lo < hi) {
                const uint32_t mid = lo + ((hi - lo) >> 1);
                const uint32_t mine_le = staged ? count_le(sruns[0], mid) + count_le(sruns[1], mid) : count_le(runs[0], mid) + count_le(runs[1], mid);
                const uint64_t c = warp_sum_u64((uint64_t)mine_le);
                if (c >= i0 + 1) hi = mid; else lo = mid + 1;
            }
            const uint32_t v0 = lo;
            uint32_t v1 = v0;
            if (pair) {
                const uint32_t c0 = staged ? count_le(sruns[0], v0) : count_le(runs[0], v0);
                const uint32_t c1 = staged ? count_le(sruns[1], v0) : count_le(runs[1], v0);
                const uint64_t n_le = warp_sum_u64((uint64_t)c0 + c1);
                if (n_le < i0 + 2) {  // element i0 + 1 is the smallest value above v0
                    uint32_t best = 0xFFFFFFFFu;
                    if (c0 < runs[0].n) best = staged ? sruns[0].p[c0] : (uint32_t)__ldg(runs[0].p + c0);
                    if (c1 < runs[1].n) { const uint32_t x = staged ? sruns[1].p[c1] : (uint32_t)__ldg(runs[1].p + c1); best = x < best ? x : best; }
                    v1 = __reduce_min_sync(0xffffffffu, best);
                }
            }
            pct[q] = pair ? __dmul_rn(__dadd_rn((double)v0, (double)v1), 0.5) : (double)v0;
        }
    }
    if (lane == 0) {
        p.emit[key] = 1;
        key_epilogue(p, key, cnt, sum, vmin, vmax, avg, sd, pct);
    }
}

}  // namespace

apm_status apm_launch_clear_slots(apm_ctx* c, uint64_t slot_mask) {
    ProfScope prof_scope_(c, APM_PG_CLEAR);
    if (!slot_mask) return APM_OK;
    unsigned gx = (c->K + 255) / 256;
    if (gx > (unsigned)c->num_sms * 4) gx = (unsigned)c->num_sms * 4;
    clear_slots_kernel<<<dim3(gx, c->S), 256, 0, c->s_main>>>(c->acc, c->slot_meta, c->K, slot_mask);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

apm_status apm_launch_tick_stats(apm_ctx* c, const WindowSlots& w, int64_t tick_ts, uint32_t tick_slot, int buf) {
    ProfScope prof_scope_(c, APM_PG_TICKSTATS);
    TickParams p;
    p.acc = c->acc; p.seen = c->seen; p.logs = c->d_log_ptrs; p.slot_meta = c->slot_meta; p.K = c->K;
    p.win = w;
    p.npct = c->cfg.num_percentiles;
    for (int i = 0; i < APM_MAX_PCT; ++i) p.pfrac[i] = c->pfrac[i];
    p.zpct[0] = c->cfg.z_series_pct[0]; p.zpct[1] = c->cfg.z_series_pct[1];
    p.tpm_div = c->tpm_div;
    p.tick_ts = tick_ts;
    p.stats = (c->cfg.flags & APM_F_KEEP_STATS) ? c->d_stats + (size_t)tick_slot * c->K : nullptr;
    p.zin = c->zin[buf];
    p.emit = c->emit[buf];
    p.unseen_records = &c->d_counters->unseen_records;
    p.with_pct = (c->cfg.flags & APM_F_PERCENTILES) && c->cfg.num_percentiles > 0;
    p.log_cap = c->log_cap;
    if (p.with_pct && w.n >= 4)  // below that a warp per key is mostly idle lanes
        tick_stats_warp_kernel<<<(c->K + 3) / 4, 128, 0, c->s_main>>>(p);
    else
        tick_stats_kernel<<<(c->K + 127) / 128, 128, 0, c->s_main>>>(p);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k2_tick() {
    const void* fns[] = {(const void*)clear_slots_kernel, (const void*)tick_stats_kernel, (const void*)tick_stats_warp_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
