// K3a: hand-written LSD radix sort of one bucket's value log.
//
// The log holds (key << 32 | elapsed) words in arrival order; after the sort every key's
// elapsed values form one ascending run, which is the array binaryConcat builds per window
// (stream_calc_stats.js:176 -> util_methods.js:84-106) without its O(W^2) splices.
// Only the significant bits are sorted: the low bits(max elapsed) of the value field and the
// low bits(K-1) of the key field, 8 bits per pass, stable.
#include "ctx.h"

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 words = 32 KB staged in shared memory

__device__ __forceinline__ uint32_t digit_of(uint64_t x, int shift) { return (uint32_t)(x >> shift) & 0xFFu; }

// lanes of the warp holding the same 8-bit digit (8 ballots)
__device__ __forceinline__ uint32_t match_digit(uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const uint32_t bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? bal : ~bal;
    }
    return peers;
}

// ---- per-tile digit histogram -> hist[digit][tile] --------------------------------------------
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint64_t* __restrict__ src, uint64_t n, int shift,
                                                             uint32_t* __restrict__ hist, uint32_t ntiles) {
    __shared__ uint32_t s_hist[RS_WARPS][256];
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&s_hist[0][0])[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint64_t base = (uint64_t)blockIdx.x * RS_TILE + (uint64_t)w * (32 * RS_ITEMS);
#pragma unroll 4
    for (int i = 0; i < RS_ITEMS; ++i) {
        const uint64_t idx = base + (uint64_t)i * 32 + lane;
        const bool ok = idx < n;
        const uint32_t d = ok ? digit_of(__ldg(src + idx), shift) : 0xFFFFFFFFu;
        // warp-aggregate equal digits so a skewed digit costs one shared atomic, not 32
        const uint32_t peers = match_digit(d & 0xFFu) & __ballot_sync(0xffffffffu, ok);
        if (ok && lane == (__ffs(peers) - 1)) atomicAdd(&s_hist[w][d], __popc(peers));
    }
    __syncthreads();
    uint32_t t = 0;
#pragma unroll
    for (int j = 0; j < RS_WARPS; ++j) t += s_hist[j][threadIdx.x];
    hist[(size_t)threadIdx.x * ntiles + blockIdx.x] = t;
}

// ---- exclusive scan of every digit row over the tiles; row totals ------------------------------
__global__ void __launch_bounds__(256) rs_scan_rows_kernel(uint32_t* __restrict__ hist, uint32_t ntiles,
                                                           uint32_t* __restrict__ totals) {
    __shared__ uint32_t s_warp[8];
    uint32_t* row = hist + (size_t)blockIdx.x * ntiles;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < ntiles; base += 256) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t x = i < ntiles ? row[i] : 0;
        uint32_t v = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += y;
        }
        if (lane == 31) s_warp[w] = v;
        __syncthreads();
        uint32_t woff = 0, tot = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (j < w) woff += s_warp[j];
            tot += s_warp[j];
        }
        if (i < ntiles) row[i] = carry + woff + v - x;
        carry += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

// ---- stable scatter of one tile ------------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(const uint64_t* __restrict__ src, uint64_t* __restrict__ dst,
                                                                uint64_t n, int shift, const uint32_t* __restrict__ hist,
                                                                const uint32_t* __restrict__ totals, uint32_t ntiles) {
    __shared__ uint64_t s_items[RS_TILE];
    __shared__ uint32_t s_cnt[RS_WARPS][256];  // per-warp digit counts, then per-warp digit offsets
    __shared__ uint32_t s_start[256];          // first local position of each digit
    __shared__ uint32_t s_gbase[256];          // global position of the tile's first item of each digit
    __shared__ uint32_t s_scan[RS_WARPS];

    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&s_cnt[0][0])[i] = 0;
    __syncthreads();

    const uint64_t tile_base = (uint64_t)blockIdx.x * RS_TILE;
    const uint64_t base = tile_base + (uint64_t)w * (32 * RS_ITEMS);
    uint64_t item[RS_ITEMS];
    uint32_t rank[RS_ITEMS];
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const uint64_t idx = base + (uint64_t)i * 32 + lane;
        item[i] = idx < n ? __ldg(src + idx) : ~0ull;
    }
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const uint64_t idx = base + (uint64_t)i * 32 + lane;
        const bool ok = idx < n;
        // out-of-range lanes take a private pseudo digit so they never join a peer group
        const uint32_t d = digit_of(item[i], shift);
        const uint32_t okmask = __ballot_sync(0xffffffffu, ok);
        const uint32_t peers = match_digit(d) & okmask;
        uint32_t before = 0;
        if (ok) before = s_cnt[w][d];
        __syncwarp();
        if (ok && lane == (__ffs(peers) - 1)) s_cnt[w][d] = before + __popc(peers);
        __syncwarp();
        rank[i] = before + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();

    // thread d owns digit d: offsets of each warp inside the digit, digit total
    {
        const int d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int j = 0; j < RS_WARPS; ++j) {
            const uint32_t c = s_cnt[j][d];
            s_cnt[j][d] = run;
            run += c;
        }
        // exclusive scan of the digit totals over the block -> s_start
        uint32_t v = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += y;
        }
        if (lane == 31) s_scan[w] = v;
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int j = 0; j < RS_WARPS; ++j)
            if (j < w) woff += s_scan[j];
        s_start[d] = woff + v - run;
        // global base of digit d = sum of totals of smaller digits + this tile's row prefix
        uint32_t tv = totals[d];
        uint32_t tincl = tv;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, tincl, o);
            if (lane >= o) tincl += y;
        }
        __syncthreads();
        if (lane == 31) s_scan[w] = tincl;
        __syncthreads();
        uint32_t toff = 0;
#pragma unroll
        for (int j = 0; j < RS_WARPS; ++j)
            if (j < w) toff += s_scan[j];
        s_gbase[d] = toff + tincl - tv + hist[(size_t)d * ntiles + blockIdx.x];
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const uint64_t idx = base + (uint64_t)i * 32 + lane;
        if (idx < n) {
            const uint32_t d = digit_of(item[i], shift);
            s_items[s_start[d] + s_cnt[w][d] + rank[i]] = item[i];
        }
    }
    __syncthreads();

    const uint32_t in_tile = (uint32_t)((n - tile_base) < (uint64_t)RS_TILE ? (n - tile_base) : RS_TILE);
    for (uint32_t j = threadIdx.x; j < in_tile; j += RS_THREADS) {
        const uint64_t x = s_items[j];
        const uint32_t d = digit_of(x, shift);
        dst[(uint64_t)s_gbase[d] + (j - s_start[d])] = x;
    }
}

// ---- accumulator planes of one bucket slot from its sorted log ----------------------------------
// After the sort every key's values are one ascending run, so count / sum / sum of squares are
// a segmented reduction over contiguous words and min / max are the run's first and last
// value: no contended atomics at all.  Each thread walks 8 consecutive words; a warp whose
// 256 words all belong to one key (the Zipf-hot keys) combines with shuffles and emits once.
constexpr int AS_THREADS = 256;
constexpr int AS_ITEMS = 8;

struct Partial {
    uint32_t cnt;
    unsigned long long sum, sq_lo, sq_hi;
};

__device__ __forceinline__ void as_emit(const AccSoA& acc, size_t base, uint32_t key, const Partial& a) {
    const size_t ai = base + key;
    atomicAdd(&acc.cnt[ai], a.cnt);
    atomicAdd(&acc.sum[ai], a.sum);
    atomicAdd(&acc.sq_lo[ai], a.sq_lo);
    if (a.sq_hi) atomicAdd(&acc.sq_hi[ai], a.sq_hi);
}

__global__ void __launch_bounds__(AS_THREADS) acc_from_sorted_kernel(const uint64_t* __restrict__ w, uint64_t n, const AccSoA acc,
                                                                    uint32_t K, int slot) {
    const size_t base = (size_t)slot * K;
    const uint64_t i0 = ((uint64_t)blockIdx.x * AS_THREADS + threadIdx.x) * AS_ITEMS;
    const bool any = i0 < n;
    uint64_t x[AS_ITEMS];
#pragma unroll
    for (int j = 0; j < AS_ITEMS; ++j) x[j] = (i0 + j < n) ? __ldcs(w + i0 + j) : ~0ull;
    const uint32_t cnt_here = any ? (uint32_t)((n - i0) < (uint64_t)AS_ITEMS ? (n - i0) : AS_ITEMS) : 0;
    uint32_t first_key = 0xFFFFFFFFu, last_key = 0xFFFFFFFFu;
    bool single = false;
    Partial run = {0, 0, 0, 0};
    if (any) {
        first_key = (uint32_t)(x[0] >> 32);
        last_key = (uint32_t)(x[cnt_here - 1] >> 32);
        single = first_key == last_key && cnt_here == AS_ITEMS;
        // run boundaries against the neighbouring words: first value of a run is its minimum, last its maximum
        const uint32_t prev_key = i0 ? (uint32_t)(__ldg(w + i0 - 1) >> 32) : 0xFFFFFFFFu;
        const uint32_t next_key = (i0 + cnt_here < n) ? (uint32_t)(__ldg(w + i0 + cnt_here) >> 32) : 0xFFFFFFFFu;
        uint32_t pk = prev_key;
#pragma unroll
        for (int j = 0; j < AS_ITEMS; ++j) {
            if ((uint32_t)j < cnt_here) {
                const uint32_t k = (uint32_t)(x[j] >> 32), v = (uint32_t)x[j];
                const uint32_t nk = ((uint32_t)(j + 1) < cnt_here) ? (uint32_t)(x[(j + 1) % AS_ITEMS] >> 32) : next_key;
                if (k != pk || (i0 + j) == 0) {
                    acc.vmin[base + k] = v;
                    acc.run_start[base + k] = (uint32_t)(i0 + j);
                }
                if (k != nk) acc.vmax[base + k] = v;
                pk = k;
            }
        }
    }
    // hot-key fast path: the whole warp holds one key
    const uint32_t k0 = __shfl_sync(0xffffffffu, first_key, 0);
    if (__all_sync(0xffffffffu, single && first_key == k0)) {
        Partial a = {AS_ITEMS, 0, 0, 0};
#pragma unroll
        for (int j = 0; j < AS_ITEMS; ++j) {
            const unsigned long long v = (uint32_t)x[j], sq = v * v;
            a.sum += v; a.sq_lo += sq & 0xFFFFFFFFull; a.sq_hi += sq >> 32;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a.cnt += __shfl_xor_sync(0xffffffffu, a.cnt, o);
            a.sum += __shfl_xor_sync(0xffffffffu, a.sum, o);
            a.sq_lo += __shfl_xor_sync(0xffffffffu, a.sq_lo, o);
            a.sq_hi += __shfl_xor_sync(0xffffffffu, a.sq_hi, o);
        }
        if ((threadIdx.x & 31) == 0) as_emit(acc, base, k0, a);
        return;
    }
    if (!any) return;
    uint32_t cur = first_key;
#pragma unroll
    for (int j = 0; j < AS_ITEMS; ++j) {
        if ((uint32_t)j < cnt_here) {
            const uint32_t k = (uint32_t)(x[j] >> 32);
            if (k != cur) {
                as_emit(acc, base, cur, run);
                run = {0, 0, 0, 0};
                cur = k;
            }
            const unsigned long long v = (uint32_t)x[j], sq = v * v;
            run.cnt += 1; run.sum += v; run.sq_lo += sq & 0xFFFFFFFFull; run.sq_hi += sq >> 32;
        }
    }
    as_emit(acc, base, cur, run);
}

__global__ void __launch_bounds__(256) clear_planes_kernel(const AccSoA acc, uint32_t K, int slot, SlotMeta* __restrict__ meta,
                                                           uint64_t n_sorted) {
    if (blockIdx.x == 0 && threadIdx.x == 0) meta[slot].sorted = n_sorted;  // the device copy of the host's bookkeeping (checkpoints)
    const size_t base = (size_t)slot * K;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)K; i += (size_t)gridDim.x * blockDim.x) {
        acc.sum[base + i] = 0; acc.sq_lo[base + i] = 0; acc.sq_hi[base + i] = 0;
        acc.cnt[base + i] = 0; acc.vmin[base + i] = 0xFFFFFFFFu; acc.vmax[base + i] = 0; acc.run_start[base + i] = 0;
    }
}

int bits_of(uint64_t v) {
    int b = 0;
    while (v) { ++b; v >>= 1; }
    return b;
}

}  // namespace

// Stable LSD sort of n words between two buffers on `st`; only the low `vbits` of the value
// field and the low `kbits` of the key field are sorted.  hist needs 256 * tiles(n) entries,
// totals 256.  Returns the buffer that holds the sorted words (a or b).
uint64_t* apm_radix_sort_words(cudaStream_t st, uint64_t* a, uint64_t* b, uint64_t n, uint32_t* hist, uint32_t* totals,
                               int vbits, int kbits, uint64_t* launches) {
    const uint32_t ntiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
    uint64_t* src = a;
    uint64_t* dst = b;
    for (int field = 0; field < 2; ++field) {
        const int lo = field == 0 ? 0 : 32, nb = field == 0 ? vbits : kbits;
        for (int bit = 0; bit < nb; bit += 8) {
            const int shift = lo + bit;
            rs_hist_kernel<<<ntiles, RS_THREADS, 0, st>>>(src, n, shift, hist, ntiles);
            rs_scan_rows_kernel<<<256, 256, 0, st>>>(hist, ntiles, totals);
            rs_scatter_kernel<<<ntiles, RS_THREADS, 0, st>>>(src, dst, n, shift, hist, totals, ntiles);
            if (launches) *launches += 3;
            uint64_t* t = src; src = dst; dst = t;
        }
    }
    return src;
}

uint64_t apm_radix_sort_tiles(uint64_t n) { return (n + RS_TILE - 1) / RS_TILE; }

int apm_bits_of(uint64_t v) { return bits_of(v); }

namespace {
// ---- one-CTA sort for small buckets -----------------------------------------------------------
// A bucket of a few hundred or thousand transactions (the reference's own deployment: ~10^2 tx
// per 10-s bucket) is not worth five launches of the tiled radix sort: one CTA loads the words,
// runs a bitonic network over them in shared memory and stores them back in place.  The words
// are distinct-or-identical 64-bit integers, so any comparison sort gives the same array.
constexpr int SS_CAP = 4096, SS_THREADS = 1024;

__global__ void __launch_bounds__(SS_THREADS) small_sort_kernel(uint64_t* __restrict__ data, uint32_t n) {
    __shared__ uint64_t s[SS_CAP];
    uint32_t m = 2;
    while (m < n) m <<= 1;
    for (uint32_t i = threadIdx.x; i < m; i += SS_THREADS) s[i] = i < n ? data[i] : ~0ull;
    __syncthreads();
    for (uint32_t k = 2; k <= m; k <<= 1) {
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t t = threadIdx.x; t < (m >> 1); t += SS_THREADS) {
                const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), l = i | j;
                const uint64_t a = s[i], b = s[l];
                const bool up = (i & k) == 0;
                if ((a > b) == up) { s[i] = b; s[l] = a; }
            }
            __syncthreads();
        }
    }
    for (uint32_t i = threadIdx.x; i < n; i += SS_THREADS) data[i] = s[i];
}
}  // namespace


namespace {
// ---- small buckets, small key spaces: sort + accumulator rebuild (+ the tick's slot clear) in ONE launch ------------
// The reference's own deployment shape (a few hundred keys, a few hundred transactions per 10-s bucket) spent its
// tick in launch latency: slot clear, one-CTA sort, plane clear and accumulator rebuild were four launches of a few
// microseconds of work each.  Here CTA b < nslots takes bucket slot[b] - load, bitonic sort in shared memory, store,
// clear the slot's planes, rebuild them from the sorted words (one thread per run start walks its run) - and the CTAs
// behind them clear the planes of the slots the tick re-labels (removeOldBuckets, stream_calc_stats.js:103-113).
struct SmallTickParams {
    int nslots;
    int slot[APM_MAX_SLOTS];
    uint32_t n[APM_MAX_SLOTS];
    unsigned long long clear_mask;   // slots the tick evicts: handled by the CTAs with index >= nslots
    int S;
};

__global__ void __launch_bounds__(SS_THREADS) small_tick_kernel(const SmallTickParams p, uint64_t* const* __restrict__ logs, const AccSoA acc,
                                                                uint32_t K, SlotMeta* __restrict__ meta) {
    __shared__ uint64_t s[SS_CAP];
    const uint32_t tid = threadIdx.x;
    if ((int)blockIdx.x >= p.nslots) {
        // the evicted slots: every (slot, key) plane entry back to empty, the slot's log length to 0
        for (int sl = 0; sl < p.S; ++sl) {
            if (!((p.clear_mask >> sl) & 1ull)) continue;
            const size_t base = (size_t)sl * K;
            for (uint32_t i = tid + (blockIdx.x - p.nslots) * SS_THREADS; i < K; i += (gridDim.x - p.nslots) * SS_THREADS) {
                acc.sum[base + i] = 0; acc.sq_lo[base + i] = 0; acc.sq_hi[base + i] = 0;
                acc.cnt[base + i] = 0; acc.vmin[base + i] = 0xFFFFFFFFu; acc.vmax[base + i] = 0; acc.run_start[base + i] = 0;
            }
            if ((int)blockIdx.x == p.nslots && tid == 0) { meta[sl].count = 0; meta[sl].sorted = 0; meta[sl].max_value = 0; }
        }
        return;
    }
    const int slot = p.slot[blockIdx.x];
    const uint32_t n = p.n[blockIdx.x];
    uint64_t* data = logs[slot];
    uint32_t m = 2;
    while (m < n) m <<= 1;
    for (uint32_t i = tid; i < m; i += SS_THREADS) s[i] = i < n ? data[i] : ~0ull;
    // this slot's planes start empty (the rebuild below writes every key that has a run)
    const size_t base = (size_t)slot * K;
    for (uint32_t i = tid; i < K; i += SS_THREADS) {
        acc.sum[base + i] = 0; acc.sq_lo[base + i] = 0; acc.sq_hi[base + i] = 0;
        acc.cnt[base + i] = 0; acc.vmin[base + i] = 0xFFFFFFFFu; acc.vmax[base + i] = 0; acc.run_start[base + i] = 0;
    }
    __syncthreads();
    for (uint32_t k = 2; k <= m; k <<= 1) {
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t t = tid; t < (m >> 1); t += SS_THREADS) {
                const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), l = i | j;
                const uint64_t a = s[i], b = s[l];
                const bool up = (i & k) == 0;
                if ((a > b) == up) { s[i] = b; s[l] = a; }
            }
            __syncthreads();
        }
    }
    for (uint32_t i = tid; i < n; i += SS_THREADS) {
        const uint64_t x = s[i];
        data[i] = x;
        const uint32_t key = (uint32_t)(x >> 32);
        if (i == 0 || (uint32_t)(s[i - 1] >> 32) != key) {
            // first word of the key's run: its value is the minimum; walk the run for count, sum, sum of squares, maximum
            uint32_t c = 0, last = 0;
            unsigned long long sum = 0, sq_lo = 0, sq_hi = 0;
            for (uint32_t q = i; q < n && (uint32_t)(s[q] >> 32) == key; ++q) {
                const unsigned long long v = (uint32_t)s[q], sq = v * v;
                ++c; sum += v; sq_lo += sq & 0xFFFFFFFFull; sq_hi += sq >> 32;
                last = (uint32_t)v;
            }
            const size_t ai = base + key;
            acc.cnt[ai] = c; acc.sum[ai] = sum; acc.sq_lo[ai] = sq_lo; acc.sq_hi[ai] = sq_hi;
            acc.vmin[ai] = (uint32_t)x; acc.vmax[ai] = last; acc.run_start[ai] = i;
        }
    }
    if (tid == 0) meta[slot].sorted = n;
}
}  // namespace

// The small-shape tick: every bucket of `slots` (n[i] words each, all at most SS_CAP) sorted and its planes rebuilt,
// and the slots of `clear_mask` emptied, in one launch.  The caller checks apm_small_tick_ok first.
bool apm_small_tick_ok(const apm_ctx* c, uint64_t largest) {
    return !c->sort_legacy && c->small_tick && largest <= (uint64_t)SS_CAP && c->K <= 16384;
}

apm_status apm_small_tick(apm_ctx* c, const int* slots, const uint64_t* n, int nslots, uint64_t clear_mask) {
    if (nslots == 0 && !clear_mask) return APM_OK;
    ProfScope prof_scope_(c, APM_PG_SORT);
    SmallTickParams p;
    p.nslots = nslots;
    for (int i = 0; i < nslots; ++i) { p.slot[i] = slots[i]; p.n[i] = (uint32_t)n[i]; }
    p.clear_mask = clear_mask;
    p.S = c->S;
    const unsigned clear_ctas = clear_mask ? (c->K + SS_THREADS - 1) / SS_THREADS : 0;
    small_tick_kernel<<<nslots + clear_ctas, SS_THREADS, 0, c->s_main>>>(p, c->d_log_ptrs, c->acc, c->K, c->slot_meta);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

// Sorts log slot `slot` (entries were appended since its last sort).
apm_status apm_sort_slot_n(apm_ctx* c, int slot, uint64_t n, uint32_t max_value) {
    ProfScope prof_scope_(c, APM_PG_SORT);
    if (n < 2) return APM_OK;
    if (n > 0xFFFFFFFFull) return apm_fail(c, APM_E_CAPACITY, "value log above 2^32 entries");
    if (apm_radix_sort_tiles(n) > c->sort_tiles_cap) return apm_fail(c, APM_E_CAPACITY, "sort scratch too small");
    if (n <= (uint64_t)SS_CAP && !c->sort_legacy) {
        small_sort_kernel<<<1, SS_THREADS, 0, c->s_main>>>(c->log_buf[slot], (uint32_t)n);
        c->launches += 1;
        CUDA_TRY(c, cudaGetLastError());
        return APM_OK;
    }
    const int vbits = bits_of(max_value), kbits = bits_of(c->K > 1 ? c->K - 1 : 0);
    uint64_t* res = nullptr;
    const bool planned = c->sort_group && c->gs_scratch && !c->sort_legacy;
    if (planned) {
        // one partition pass + one pass of shared-memory sorts, planned from the last sorted bucket
        cudaError_t e = cudaSuccess;
        if (apm_groupsort(c->s_main, &c->gs, c->gs_scratch, c->K, c->log_buf[slot], c->log_buf[c->S], n, vbits, &c->launches, &e))
            res = c->log_buf[c->S];
        CUDA_TRY(c, e);
    }
    const bool by_plan = res != nullptr;
    if (!res && !c->sort_legacy && c->sort_scratch) {
        cudaError_t e = cudaSuccess;
        res = apm_onesweep_sort(c->s_main, c->log_buf[slot], c->log_buf[c->S], n, c->sort_scratch, c->sort_scratch_bytes, vbits,
                                kbits, &c->launches, &e);
        CUDA_TRY(c, e);
    }
    if (!res)
        res = apm_radix_sort_words(c->s_main, c->log_buf[slot], c->log_buf[c->S], n, c->sort_hist, c->sort_base, vbits, kbits,
                                   &c->launches);
    CUDA_TRY(c, cudaGetLastError());
    if (planned && !by_plan) CUDA_TRY(c, apm_groupsort_plan_from(c->s_main, &c->gs, c->gs_scratch, c->K, res, n, &c->launches));
    if (res != c->log_buf[slot]) {  // result sits in the spare buffer: swap roles
        c->log_buf[c->S] = c->log_buf[slot];
        c->log_buf[slot] = res;
        CUDA_TRY(c, cudaMemcpyAsync(c->d_log_ptrs + slot, &c->log_buf[slot], sizeof(uint64_t*), cudaMemcpyHostToDevice,
                                    c->s_main));
    }
    return APM_OK;
}

// Rebuilds the accumulator planes of `slot` from its sorted log of n words.
apm_status apm_acc_from_sorted(apm_ctx* c, int slot, uint64_t n) {
    ProfScope prof_scope_(c, APM_PG_REDUCE);
    unsigned g = (c->K + 255) / 256;
    clear_planes_kernel<<<g < (unsigned)c->num_sms * 4 ? g : (unsigned)c->num_sms * 4, 256, 0, c->s_main>>>(c->acc, c->K, slot, c->slot_meta, n);
    c->launches += 1;
    if (n) {
        const uint64_t per_block = (uint64_t)AS_THREADS * AS_ITEMS;
        acc_from_sorted_kernel<<<(unsigned)((n + per_block - 1) / per_block), AS_THREADS, 0, c->s_main>>>(c->log_buf[slot], n, c->acc,
                                                                                                      c->K, slot);
        c->launches += 1;
    }
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k3_sort() {
    const void* fns[] = {(const void*)small_tick_kernel, (const void*)acc_from_sorted_kernel, (const void*)clear_planes_kernel, (const void*)rs_hist_kernel, (const void*)rs_scan_rows_kernel, (const void*)rs_scatter_kernel, (const void*)small_sort_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
