// K3g: the bucket sort as ONE partition pass and ONE pass of shared-memory sorts.
//
// Same job as k3_onesweep.cu - order a bucket's value log by (key, elapsed) so every key's durations form one
// ascending run (what binaryConcat builds, util_methods.js:84-106) - but the five global passes of an LSD radix sort
// (each a load -> rank -> publish -> look back -> scatter chain across one wave of tiles) are replaced by two:
//
//   plan     every G-th word of the PREVIOUS sorted bucket is a splitter.  Traffic is stationary from one bucket to the
//            next, so the splitters cut the new bucket into groups of about G words whatever its skew: a Zipf-hot key
//            is cut by VALUE into many groups, a thousand tail keys share one.  A word equal to a splitter goes to that
//            splitter's own "equal" group, which needs no sorting however large it is (a hot key has a few thousand
//            distinct durations among its 10^5 words).
//   count    one read of the log: words per group (shared-memory histogram per CTA).
//   scatter  every tile of 4096 words reserves room in each group with one atomic per (tile, group) and writes its
//            words there.  No order is kept inside a group, so no tile waits for another: there is no look-back chain.
//   sort     one CTA per group: load (<= 4096 words), LSD radix in shared memory over the bits that differ inside the
//            group only (a hot key's value slice: one pass), write back in place, and hand the next plan its splitters.
//
// Correctness does not depend on the plan: groups are ranges of the (key, value) order, so sorted groups in group
// order ARE the sorted bucket.  A plan that fits badly (first bucket of a context, a shift in the traffic) only makes
// groups too large for shared memory; those are sorted by their CTA through global memory, and a bucket without any
// plan takes the onesweep path and leaves a plan behind.
//
// Status: correct on every test, NOT the default (APM_SORT_GROUP=1 turns it on).  Measured on B200 for cfg3's bucket
// (1.67 M words, 10 k keys, Zipf; scripts/micro_sort, profiles/r2_ncu_groupsort_summary.txt): 103 us against 106 us for
// the four radix passes.  The chain is gone, but what replaced it costs as much: count 25 us + scatter 37 us - both
// dominated by the group lookup, a divergent search through the hot keys' splitters and one 32-line gather per warp
// for the table, 8.2 M and 6.4 M warp instructions - and sort 35 us (13.8 M warp instructions: the ballot ranking of the
// shared-memory passes costs ~80 instructions per word and pass, the same as a global pass).
#include <cstdlib>

#include "ctx.h"

namespace {

constexpr int GS_THREADS = 256;
constexpr int GS_WARPS = GS_THREADS / 32;
constexpr int GS_ITEMS = 16;
constexpr int GS_TILE = GS_THREADS * GS_ITEMS;  // 4096 words
constexpr int GS_GMAX = GS_TILE;                // a group of up to this many words is sorted in shared memory
constexpr uint32_t GS_TARGET = 2048;            // words per group the plan aims at: room for drift before GS_GMAX

// group of a word: 2c for a word strictly between splitters c-1 and c, 2c - 1 for a word equal to splitter c-1
// (c = number of splitters <= w).  tbl[k] = number of splitters whose key field is below k, so only the splitters
// that carry the word's own key - none for almost every key - are searched.
__device__ __forceinline__ uint32_t gs_classify(uint64_t w, const uint64_t* __restrict__ split, const uint32_t* __restrict__ tbl) {
    const uint32_t k = (uint32_t)(w >> 32);
    const uint32_t first = __ldg(tbl + k);
    uint32_t lo = first, hi = __ldg(tbl + k + 1);
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(split + mid) <= w) lo = mid + 1; else hi = mid;
    }
    return (lo > first && __ldg(split + lo - 1) == w) ? 2u * lo - 1u : 2u * lo;
}

// ---- count: words per group --------------------------------------------------------------------------------
__global__ void __launch_bounds__(GS_THREADS) gs_count_kernel(const uint64_t* __restrict__ src, uint32_t n, const uint64_t* __restrict__ split,
                                                              const uint32_t* __restrict__ tbl, uint32_t ngroups, uint32_t* __restrict__ ghist) {
    extern __shared__ uint32_t gs_smem_hist[];
    for (uint32_t i = threadIdx.x; i < ngroups; i += GS_THREADS) gs_smem_hist[i] = 0;
    __syncthreads();
    // warp-uniform trip counts, and the lanes rejoin after every lookup: a lane that searches a hot key's splitters must
    // not leave its warp split for the rest of the loop
    const uint32_t tiles = (n + GS_TILE - 1) / GS_TILE;
    for (uint32_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const uint32_t base = tile * GS_TILE;
        uint64_t x[4];
        for (int it = 0; it < GS_ITEMS; it += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t idx = base + (it + u) * GS_THREADS + threadIdx.x;
                x[u] = idx < n ? __ldg(src + idx) : 0ull;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t idx = base + (it + u) * GS_THREADS + threadIdx.x;
                if (idx < n) atomicAdd(&gs_smem_hist[gs_classify(x[u], split, tbl)], 1u);
                __syncwarp();
            }
        }
    }
    __syncthreads();
    for (uint32_t g = threadIdx.x; g < ngroups; g += GS_THREADS) {
        const uint32_t v = gs_smem_hist[g];
        if (v) atomicAdd(&ghist[g], v);
    }
}

// ---- exclusive scan of the group counts (one CTA; at most 4097 groups) -> gstart[0 .. ngroups], cursors zeroed --------
__global__ void __launch_bounds__(1024) gs_scan_kernel(const uint32_t* __restrict__ ghist, uint32_t ngroups, uint32_t* __restrict__ gstart,
                                                       uint32_t* __restrict__ cursor) {
    __shared__ uint32_t s_warp[32];
    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    constexpr int PER = 5;  // 1024 x 5 >= 4097 + 1
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const uint32_t g = tid * PER + j;
        v[j] = g < ngroups ? ghist[g] : 0u;
        sum += v[j];
    }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += y;
    }
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    if (w == 0) {
        uint32_t x = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= (uint32_t)o) x += y;
        }
        s_warp[lane] = x;
    }
    __syncthreads();
    uint32_t run = incl - sum + (w ? s_warp[w - 1] : 0u);
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const uint32_t g = tid * PER + j;
        if (g <= ngroups) gstart[g] = run;   // gstart[ngroups] = n
        if (g < ngroups) cursor[g] = 0;
        run += v[j];
    }
}

// ---- scatter: every word to its group, unordered inside the group -------------------------------------------------
__global__ void __launch_bounds__(GS_THREADS) gs_scatter_kernel(const uint64_t* __restrict__ src, uint64_t* __restrict__ dst, uint32_t n,
                                                                const uint64_t* __restrict__ split, const uint32_t* __restrict__ tbl,
                                                                uint32_t ngroups, const uint32_t* __restrict__ gstart,
                                                                uint32_t* __restrict__ cursor) {
    extern __shared__ uint32_t gs_smem_cnt[];  // [ngroups]: the tile's words per group, then where they go
    for (uint32_t i = threadIdx.x; i < ngroups; i += GS_THREADS) gs_smem_cnt[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * GS_TILE;
    uint64_t w[GS_ITEMS];
    uint32_t g[GS_ITEMS];
    uint16_t r[GS_ITEMS];
#pragma unroll
    for (int i = 0; i < GS_ITEMS; ++i) {
        const uint32_t idx = base + i * GS_THREADS + threadIdx.x;
        w[i] = idx < n ? __ldcs(src + idx) : 0ull;
    }
#pragma unroll
    for (int i = 0; i < GS_ITEMS; ++i) {
        const uint32_t idx = base + i * GS_THREADS + threadIdx.x;
        g[i] = 0;
        r[i] = 0;
        if (idx < n) {
            g[i] = gs_classify(w[i], split, tbl);
            r[i] = (uint16_t)atomicAdd(&gs_smem_cnt[g[i]], 1u);  // any order will do: the group is sorted afterwards
        }
        __syncwarp();  // lanes that searched a hot key's splitters rejoin the others
    }
    __syncthreads();
    for (uint32_t q = threadIdx.x; q < ngroups; q += GS_THREADS) {
        const uint32_t c = gs_smem_cnt[q];
        if (c) gs_smem_cnt[q] = __ldg(gstart + q) + atomicAdd(&cursor[q], c);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < GS_ITEMS; ++i) {
        const uint32_t idx = base + i * GS_THREADS + threadIdx.x;
        if (idx < n) dst[gs_smem_cnt[g[i]] + r[i]] = w[i];
    }
}

// ---- sort: one CTA per group -----------------------------------------------------------------------------------
// logical word: the low vbits of the duration with the key packed directly above - the bits between are always zero
// and would only cost radix passes
__device__ __forceinline__ uint64_t gs_logical(uint64_t x, int vbits) {
    return vbits >= 32 ? x : ((x & ((1ull << vbits) - 1ull)) | ((x >> 32) << vbits));
}
__device__ __forceinline__ uint64_t gs_physical(uint64_t l, int vbits) {
    return vbits >= 32 ? l : ((l & ((1ull << vbits) - 1ull)) | ((l >> vbits) << 32));
}

struct GsSortSmem {
    uint64_t a[GS_GMAX];
    uint64_t b[GS_GMAX];
    uint16_t wcnt[GS_WARPS][256];  // words per (warp, digit), then the warp's offset inside the digit
    uint32_t start[256];           // first position of each digit
    uint32_t run[256];             // slow path: words of each digit written by the tiles before this one
    uint32_t scan[GS_WARPS];
    unsigned long long diff;       // OR of (word ^ first word): the bits that differ inside the group
};

// Stable rank of up to ipt x 256 items by one 8-bit digit.  Item i of a thread sits at index w * 32 * ipt + i * 32 +
// lane (index order == (warp, i, lane) order, so ranks taken in that order are stable).  Returns the per-item rank
// inside (warp, digit); sm.wcnt ends as each warp's offset inside its digit and sm.start as the tile's exclusive scan
// over the digits (the count of a digit is left in `count_out` of the thread that owns it).
template <typename Load>
__device__ __forceinline__ void gs_rank_tile(GsSortSmem& sm, int ipt, uint32_t m_tile, int shift, Load item_at, uint16_t (&rank)[GS_ITEMS],
                                             uint32_t& count_out) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (uint32_t i = tid; i < (uint32_t)(GS_WARPS * 256 / 2); i += GS_THREADS) reinterpret_cast<uint32_t*>(&sm.wcnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t lanes_below = (1u << lane) - 1u;
    uint16_t* my = sm.wcnt[w];
#pragma unroll
    for (int i = 0; i < GS_ITEMS; ++i) {
        if (i >= ipt) break;  // uniform over the CTA
        const uint32_t idx = w * 32u * (uint32_t)ipt + (uint32_t)i * 32u + lane;
        const bool ok = idx < m_tile;
        const uint32_t d = ok ? (uint32_t)(item_at(idx) >> shift) & 0xFFu : 0xFFu;  // padding ranks behind every word of its warp
        uint32_t peers = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            peers &= bit ? bal : ~bal;
        }
        const uint32_t before = my[d];
        __syncwarp();
        if ((peers & lanes_below) == 0) my[d] = (uint16_t)(before + __popc(peers));
        __syncwarp();
        rank[i] = (uint16_t)(before + __popc(peers & lanes_below));
    }
    __syncthreads();
    // thread d owns digit d: the warps' offsets inside it, then the exclusive scan over the digits
    {
        const uint32_t d = tid;
        uint32_t run = 0;
#pragma unroll
        for (int j = 0; j < GS_WARPS; ++j) {
            const uint32_t c = sm.wcnt[j][d];
            sm.wcnt[j][d] = (uint16_t)run;
            run += c;
        }
        count_out = run;
        uint32_t v = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= (uint32_t)o) v += y;
        }
        if (lane == 31) sm.scan[w] = v;
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int j = 0; j < GS_WARPS; ++j)
            if ((uint32_t)j < w) woff += sm.scan[j];
        sm.start[d] = woff + v - run;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(GS_THREADS, 3) gs_sort_kernel(uint64_t* __restrict__ data, uint64_t* __restrict__ scratch,
                                                                const uint32_t* __restrict__ gstart, int vbits, uint32_t sample_every,
                                                                uint32_t num_samples, uint64_t* __restrict__ split_next) {
    extern __shared__ __align__(16) unsigned char gs_sort_raw[];
    GsSortSmem& sm = *reinterpret_cast<GsSortSmem*>(gs_sort_raw);
    const uint32_t g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t start = __ldg(gstart + g), m = __ldg(gstart + g + 1) - start;
    if (m == 0) return;
    uint64_t* grp = data + start;
    const bool equal_group = (g & 1u) != 0;  // every word equals one splitter: already in order
    if (!equal_group && m > 1) {
        if (tid == 0) sm.diff = 0ull;
        __syncthreads();
        if (m <= (uint32_t)GS_GMAX) {
            // ---- in shared memory
            const uint64_t first = gs_logical(grp[0], vbits);
            unsigned long long diff = 0;
            for (uint32_t i = tid; i < m; i += GS_THREADS) {
                const uint64_t l = gs_logical(__ldcs(grp + i), vbits);
                sm.a[i] = l;
                diff |= l ^ first;
            }
            diff = __reduce_or_sync(0xffffffffu, (uint32_t)diff) | ((unsigned long long)__reduce_or_sync(0xffffffffu, (uint32_t)(diff >> 32)) << 32);
            if (lane == 0 && diff) atomicOr(&sm.diff, diff);
            __syncthreads();
            const unsigned long long dbits = sm.diff;
            const int nbits = dbits ? 64 - __clzll((long long)dbits) : 0;
            const int ipt = (int)((m + GS_THREADS - 1) / GS_THREADS);
            uint64_t* a = sm.a;
            uint64_t* b = sm.b;
            for (int shift = 0; shift < nbits; shift += 8) {
                uint16_t rank[GS_ITEMS];
                uint32_t cnt_unused;
                gs_rank_tile(sm, ipt, m, shift, [&](uint32_t idx) { return a[idx]; }, rank, cnt_unused);
#pragma unroll
                for (int i = 0; i < GS_ITEMS; ++i) {
                    const uint32_t idx = w * 32u * (uint32_t)ipt + (uint32_t)i * 32u + lane;
                    if (i < ipt && idx < m) {
                        const uint64_t x = a[idx];
                        const uint32_t d = (uint32_t)(x >> shift) & 0xFFu;
                        b[sm.start[d] + sm.wcnt[w][d] + rank[i]] = x;
                    }
                }
                __syncthreads();
                uint64_t* t = a; a = b; b = t;
            }
            for (uint32_t i = tid; i < m; i += GS_THREADS) grp[i] = gs_physical(a[i], vbits);
        } else {
            // ---- a group the plan did not foresee: the same radix passes, tile by tile through global memory, between
            //      the group's place in `data` and the same place in `scratch` (the unsorted log, dead by now)
            const uint64_t first = gs_logical(grp[0], vbits);
            unsigned long long diff = 0;
            for (uint32_t i = tid; i < m; i += GS_THREADS) diff |= gs_logical(grp[i], vbits) ^ first;
            diff = __reduce_or_sync(0xffffffffu, (uint32_t)diff) | ((unsigned long long)__reduce_or_sync(0xffffffffu, (uint32_t)(diff >> 32)) << 32);
            if (lane == 0 && diff) atomicOr(&sm.diff, diff);
            __syncthreads();
            const unsigned long long dbits = sm.diff;
            const int nbits = dbits ? 64 - __clzll((long long)dbits) : 0;
            uint64_t* src = grp;
            uint64_t* dst = scratch + start;
            for (int shift = 0; shift < nbits; shift += 8) {
                // digit totals of the whole group -> where each digit starts
                sm.run[tid] = 0;
                __syncthreads();
                for (uint32_t i = tid; i < m; i += GS_THREADS) atomicAdd(&sm.run[(uint32_t)(gs_logical(src[i], vbits) >> shift) & 0xFFu], 1u);
                __syncthreads();
                {
                    const uint32_t c = sm.run[tid];
                    uint32_t v = c;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
                        if (lane >= (uint32_t)o) v += y;
                    }
                    if (lane == 31) sm.scan[w] = v;
                    __syncthreads();
                    uint32_t woff = 0;
#pragma unroll
                    for (int j = 0; j < GS_WARPS; ++j)
                        if ((uint32_t)j < w) woff += sm.scan[j];
                    __syncthreads();
                    sm.run[tid] = woff + v - c;
                }
                __syncthreads();
                for (uint32_t t0 = 0; t0 < m; t0 += GS_TILE) {
                    const uint32_t mt = m - t0 < (uint32_t)GS_TILE ? m - t0 : (uint32_t)GS_TILE;
                    const uint64_t* tile = src + t0;
                    uint16_t rank[GS_ITEMS];
                    uint32_t my_count;
                    gs_rank_tile(sm, GS_ITEMS, mt, shift, [&](uint32_t idx) { return gs_logical(tile[idx], vbits); }, rank, my_count);
#pragma unroll
                    for (int i = 0; i < GS_ITEMS; ++i) {
                        const uint32_t idx = w * 32u * GS_ITEMS + (uint32_t)i * 32u + lane;
                        if (idx < mt) {
                            const uint64_t x = tile[idx];
                            const uint32_t d = (uint32_t)(gs_logical(x, vbits) >> shift) & 0xFFu;
                            dst[sm.run[d] + sm.wcnt[w][d] + rank[i]] = x;
                        }
                    }
                    __syncthreads();
                    // thread d counted its digit over the whole tile, padding included (digit 255 only)
                    uint32_t real = my_count;
                    if (tid == 255) real -= (uint32_t)GS_TILE - mt;
                    sm.run[tid] += real;
                    __threadfence_block();
                    __syncthreads();
                }
                uint64_t* t = src; src = dst; dst = t;
            }
            if (src != grp)
                for (uint32_t i = tid; i < m; i += GS_THREADS) grp[i] = src[i];
        }
        __threadfence_block();
        __syncthreads();
    }
    // the next plan's splitters: every sample_every-th word of the sorted bucket, picked by whoever holds it
    if (split_next && sample_every) {
        const uint32_t j0 = start ? (start + sample_every - 1) / sample_every - 1 : 0;  // first j with (j + 1) * sample_every >= start
        for (uint32_t j = j0 + tid; j < num_samples; j += GS_THREADS) {
            const uint64_t pos = (uint64_t)(j + 1) * sample_every;
            if (pos >= (uint64_t)start + m) break;
            if (pos >= start) split_next[j] = grp[pos - start];
        }
    }
}

// ---- plan from a bucket that some other sort put in order ----------------------------------------------------------
__global__ void __launch_bounds__(256) gs_sample_kernel(const uint64_t* __restrict__ sorted, uint32_t sample_every, uint32_t num_samples,
                                                        uint64_t* __restrict__ split) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < num_samples) split[j] = sorted[(uint64_t)(j + 1) * sample_every];
}

// tbl[k] = splitters whose key field is below k, k = 0 .. K
__global__ void __launch_bounds__(256) gs_table_kernel(const uint64_t* __restrict__ split, uint32_t P, uint32_t K, uint32_t* __restrict__ tbl) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > K) return;
    uint32_t lo = 0, hi = P;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if ((uint32_t)(split[mid] >> 32) < k) lo = mid + 1; else hi = mid;
    }
    tbl[k] = lo;
}

}  // namespace

size_t apm_groupsort_scratch_bytes(uint32_t K) {
    // [split x 2][ghist][gstart][cursor][tbl x 2]
    const size_t groups = 2 * (size_t)APM_GS_MAX_SPLIT + 2;
    size_t b = 2 * (size_t)APM_GS_MAX_SPLIT * sizeof(uint64_t);
    b += 3 * groups * sizeof(uint32_t);
    b += 2 * ((size_t)K + 2) * sizeof(uint32_t);
    return b + 256;
}

namespace {
struct GsLayout {
    uint64_t* split[2];
    uint32_t* ghist;
    uint32_t* gstart;
    uint32_t* cursor;
    uint32_t* tbl[2];
};
GsLayout gs_layout(unsigned char* scratch, uint32_t K) {
    const size_t groups = 2 * (size_t)APM_GS_MAX_SPLIT + 2;
    GsLayout l;
    l.split[0] = reinterpret_cast<uint64_t*>(scratch);
    l.split[1] = l.split[0] + APM_GS_MAX_SPLIT;
    l.ghist = reinterpret_cast<uint32_t*>(l.split[1] + APM_GS_MAX_SPLIT);
    l.gstart = l.ghist + groups;
    l.cursor = l.gstart + groups;
    l.tbl[0] = l.cursor + groups;
    l.tbl[1] = l.tbl[0] + (size_t)K + 2;
    return l;
}

// splitters a plan built from n sorted words holds, and their spacing (0: no plan - the bucket is too small or too large)
void gs_plan_shape(uint64_t n, uint32_t* sample_every, uint32_t* num_samples) {
    *sample_every = 0;
    *num_samples = 0;
    if (n < 2ull * GS_TARGET || n > (uint64_t)APM_GS_MAX_SPLIT * GS_TARGET) return;
    *sample_every = GS_TARGET;
    *num_samples = (uint32_t)((n - 1) / GS_TARGET);
    if (*num_samples > APM_GS_MAX_SPLIT) *num_samples = APM_GS_MAX_SPLIT;
}
}  // namespace

// Builds the plan for the next bucket from `sorted` (n words in order, sorted by another path).
cudaError_t apm_groupsort_plan_from(cudaStream_t st, apm_groupsort_state* gs, unsigned char* scratch, uint32_t K, const uint64_t* sorted,
                                    uint64_t n, uint64_t* launches) {
    uint32_t every, P;
    gs_plan_shape(n, &every, &P);
    gs->P = 0;
    if (!P) return cudaSuccess;
    GsLayout l = gs_layout(scratch, K);
    const int nx = gs->cur ^ 1;
    gs_sample_kernel<<<(P + 255) / 256, 256, 0, st>>>(sorted, every, P, l.split[nx]);
    gs_table_kernel<<<(K + 1 + 255) / 256, 256, 0, st>>>(l.split[nx], P, K, l.tbl[nx]);
    if (launches) *launches += 2;
    gs->cur = nx;
    gs->P = P;
    return cudaGetLastError();
}

// Sorts n words from `src` into `dst` with the current plan and leaves the plan for the next bucket behind.
// Returns false (nothing launched) when there is no plan or the bucket does not fit this path.
bool apm_groupsort(cudaStream_t st, apm_groupsort_state* gs, unsigned char* scratch, uint32_t K, uint64_t* src, uint64_t* dst, uint64_t n,
                   int vbits, uint64_t* launches, cudaError_t* err) {
    *err = cudaSuccess;
    if (!gs->P || n < 2 || n >= (1ull << 31)) return false;
    GsLayout l = gs_layout(scratch, K);
    const int cur = gs->cur, nx = cur ^ 1;
    const uint32_t P = gs->P, ngroups = 2 * P + 1;
    static unsigned long long attr_done = 0;
    if (apm_first_on_device(&attr_done)) {
        *err = cudaFuncSetAttribute(gs_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GsSortSmem));
        if (*err != cudaSuccess) { attr_done = 0; return false; }
    }
    *err = cudaMemsetAsync(l.ghist, 0, (size_t)ngroups * sizeof(uint32_t), st);
    if (*err != cudaSuccess) return false;
    const unsigned tiles = (unsigned)((n + GS_TILE - 1) / GS_TILE);
    unsigned cg = (unsigned)apm_sm_count() * 4;
    if (cg > tiles) cg = tiles;
    gs_count_kernel<<<cg, GS_THREADS, (size_t)ngroups * 4, st>>>(src, (uint32_t)n, l.split[cur], l.tbl[cur], ngroups, l.ghist);
    gs_scan_kernel<<<1, 1024, 0, st>>>(l.ghist, ngroups, l.gstart, l.cursor);
    gs_scatter_kernel<<<tiles, GS_THREADS, (size_t)ngroups * 4, st>>>(src, dst, (uint32_t)n, l.split[cur], l.tbl[cur], ngroups, l.gstart, l.cursor);
    uint32_t every, Pn;
    gs_plan_shape(n, &every, &Pn);
    gs_sort_kernel<<<ngroups, GS_THREADS, sizeof(GsSortSmem), st>>>(dst, src, l.gstart, vbits, every, Pn, Pn ? l.split[nx] : nullptr);
    if (Pn) gs_table_kernel<<<(K + 1 + 255) / 256, 256, 0, st>>>(l.split[nx], Pn, K, l.tbl[nx]);
    if (launches) *launches += Pn ? 5 : 4;
    gs->cur = nx;
    gs->P = Pn;
    *err = cudaGetLastError();
    return *err == cudaSuccess;
}

cudaError_t apm_preload_k3_groupsort() {
    const void* fns[] = {(const void*)gs_count_kernel, (const void*)gs_scan_kernel, (const void*)gs_scatter_kernel, (const void*)gs_sort_kernel,
                         (const void*)gs_sample_kernel, (const void*)gs_table_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
