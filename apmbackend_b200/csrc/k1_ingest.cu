// K0/K1: tick finder and keyed segmented reduction of a transaction segment.
//
// Replaces, for a whole batch at once, what consumeMsg does per message
// (stream_calc_stats.js:331-371): generateBucketLabel (:89-91), the
// `bucketLabel > latestBucket` test (:348) and addData (:115-134).
#include <algorithm>
#include <cstddef>
#include <cstring>

#include "ctx.h"

namespace {

constexpr int TF_THREADS = 256;
constexpr int TF_ITEMS = 16;
constexpr int TF_TILE = TF_THREADS * TF_ITEMS;

__device__ __forceinline__ long long warp_max_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        long long other = __shfl_xor_sync(0xffffffffu, v, o);
        v = other > v ? other : v;
    }
    return v;
}

// ---- K0a: per-tile maximum bucket label (8 B/tx read) -------------------------------------
__global__ void __launch_bounds__(TF_THREADS) tf_tile_max_kernel(const int64_t* __restrict__ ts, uint64_t n,
                                                                 int64_t bucket_ms, double inv_bucket,
                                                                 long long* __restrict__ tile_max) {
    __shared__ long long s_max[TF_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * TF_TILE;
    // the label is monotonic in the timestamp, so the largest label of a tile is the label of its largest timestamp:
    // the tile is reduced on the raw 64-bit values (negative = invalid = -1) and labelled once, which leaves this
    // kernel with one compare per transaction instead of the label arithmetic (it ran at a quarter of HBM speed)
    long long m = -1;
    int64_t t[TF_ITEMS];
#pragma unroll
    for (int i = 0; i < TF_ITEMS; ++i) {
        const uint64_t idx = base + (uint64_t)i * TF_THREADS + threadIdx.x;
        t[i] = idx < n ? __ldg(ts + idx) : -1;
    }
#pragma unroll
    for (int i = 0; i < TF_ITEMS; ++i) m = t[i] > m ? t[i] : m;
    m = warp_max_ll(m);
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        long long v = threadIdx.x < TF_THREADS / 32 ? s_max[threadIdx.x] : -1;
        v = warp_max_ll(v);
        if (threadIdx.x == 0) tile_max[blockIdx.x] = v >= 0 ? apm_label(v, bucket_ms, inv_bucket) : -1;
    }
}

// ---- K0b: exclusive running maximum over the tiles, carry-in = latestBucket ----------------
__global__ void __launch_bounds__(1024) tf_scan_kernel(const long long* __restrict__ tile_max, uint64_t ntiles,
                                                       long long latest, long long* __restrict__ tile_prefix) {
    __shared__ long long s_warp[32];
    const uint64_t per = (ntiles + 1023) / 1024;
    const uint64_t lo = (uint64_t)threadIdx.x * per;
    const uint64_t hi = lo + per < ntiles ? lo + per : ntiles;
    long long m = -1;
    for (uint64_t i = lo; i < hi; ++i) m = tile_max[i] > m ? tile_max[i] : m;
    // inclusive scan (max) across the block
    long long v = m;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long other = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o && other > v) v = other;
    }
    if (lane == 31) s_warp[w] = v;
    __syncthreads();
    if (w == 0) {
        long long x = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long other = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o && other > x) x = other;
        }
        s_warp[lane] = x;
    }
    __syncthreads();
    long long excl = __shfl_up_sync(0xffffffffu, v, 1);
    if (lane == 0) excl = -1;
    if (w > 0 && s_warp[w - 1] > excl) excl = s_warp[w - 1];
    long long run = excl > latest ? excl : latest;
    for (uint64_t i = lo; i < hi; ++i) {
        tile_prefix[i] = run;
        run = tile_max[i] > run ? tile_max[i] : run;
    }
}

// ---- K0c: emit the positions where the label exceeds everything before it ------------------
__global__ void __launch_bounds__(TF_THREADS) tf_emit_kernel(const int64_t* __restrict__ ts, uint64_t n,
                                                             int64_t bucket_ms, double inv_bucket,
                                                             const long long* __restrict__ tile_max,
                                                             const long long* __restrict__ tile_prefix,
                                                             TickMark* __restrict__ ticks, unsigned int* __restrict__ nticks,
                                                             unsigned int cap) {
    const long long carry = tile_prefix[blockIdx.x];
    if (tile_max[blockIdx.x] <= carry) return;  // no tick inside this tile
    __shared__ long long s_warp[TF_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * TF_TILE + (uint64_t)threadIdx.x * TF_ITEMS;
    long long lab[TF_ITEMS];
    long long m = -1;
#pragma unroll
    for (int i = 0; i < TF_ITEMS; ++i) {
        uint64_t idx = base + i;
        long long l = -1;
        if (idx < n) {
            int64_t t = ts[idx];
            if (t >= 0) l = apm_label(t, bucket_ms, inv_bucket);
        }
        lab[i] = l;
        m = l > m ? l : m;
    }
    long long v = m;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long other = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o && other > v) v = other;
    }
    if (lane == 31) s_warp[w] = v;
    __syncthreads();
    long long excl = __shfl_up_sync(0xffffffffu, v, 1);
    if (lane == 0) excl = -1;
    for (int j = 0; j < w; ++j) excl = s_warp[j] > excl ? s_warp[j] : excl;
    long long run = excl > carry ? excl : carry;
#pragma unroll
    for (int i = 0; i < TF_ITEMS; ++i) {
        if (lab[i] > run) {
            run = lab[i];
            unsigned int k = atomicAdd(nticks, 1u);
            if (k < cap) {
                ticks[k].pos = base + i;
                ticks[k].label = lab[i];
            }
        }
    }
}

// ---- K0d: how many log entries each segment adds to each bucket ---------------------------------
// With the tick positions known, a transaction's segment - and with it the latestBucket in force
// when consumeMsg would have handled it - is known too, so the outcome of addData's tests is a
// pure function of the transaction: this kernel applies exactly the tests of the ingest kernel
// below and counts, per segment s and bucket age d = latest - label, the transactions that will be
// appended to a value log, plus the largest duration among them.  The host reads the table once
// per chunk and from then on knows every bucket's log length (grid sizes, number of radix passes,
// which buckets need a sort) without asking the device at each tick.
struct SegHistParams {
    const uint32_t* svc;
    const uint32_t* jvm;
    const int64_t* ts;
    const int32_t* el;
    uint64_t n;
    int64_t bucket_ms;
    double inv_bucket;
    int64_t latest0;           // latestBucket before the chunk
    int keep;
    uint32_t num_services, num_jvms;
    const TickMark* ticks;     // sorted by position
    uint32_t nt;
    uint32_t* counts;          // [(nt + 1)][keep + 1]
    uint32_t* max_value;
};

constexpr int SH_THREADS = 256;
constexpr int SH_ITEMS = 16;
constexpr int SH_TILE = SH_THREADS * SH_ITEMS;
constexpr int SH_SPAN = 32;  // segments of one tile counted in shared memory
constexpr uint32_t SH_NONE = 0xFFFFFFFFu;

// number of tick marks with pos <= i among ticks[lo, hi)
__device__ __forceinline__ uint32_t sh_ticks_le(const TickMark* ticks, uint32_t lo, uint32_t hi, uint64_t i) {
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (ticks[mid].pos <= i) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(SH_THREADS) seg_hist_kernel(const SegHistParams p) {
    __shared__ uint32_t s_cnt[SH_SPAN * APM_MAX_SLOTS];
    __shared__ unsigned long long s_pos[SH_SPAN];
    __shared__ long long s_lab[SH_SPAN];
    __shared__ uint32_t s_seg_lo, s_nseg, s_max;
    const int tid = threadIdx.x, lane = tid & 31;
    const uint64_t tile_begin = (uint64_t)blockIdx.x * SH_TILE;
    const uint64_t tile_end = tile_begin + SH_TILE < p.n ? tile_begin + SH_TILE : p.n;
    if (tid == 0) {
        const uint32_t lo = sh_ticks_le(p.ticks, 0, p.nt, tile_begin);
        const uint32_t hi = sh_ticks_le(p.ticks, lo, p.nt, tile_end - 1);
        s_seg_lo = lo;          // segment of the tile's first transaction
        s_nseg = hi - lo + 1;   // segments the tile touches
        s_max = 0;
    }
    __syncthreads();
    const uint32_t seg_lo = s_seg_lo, nseg = s_nseg;
    const bool local = nseg <= (uint32_t)SH_SPAN;
    const int width = p.keep + 1;
    if (local) {
        for (uint32_t j = tid; j < nseg; j += SH_THREADS) {
            const uint32_t seg = seg_lo + j;
            s_lab[j] = seg == 0 ? p.latest0 : p.ticks[seg - 1].label;
            s_pos[j] = j + 1 < nseg ? p.ticks[seg].pos : ~0ull;  // where the next segment starts
        }
        for (uint32_t j = tid; j < nseg * APM_MAX_SLOTS; j += SH_THREADS) s_cnt[j] = 0;
    }
    __syncthreads();
    uint32_t my_max = 0;
#pragma unroll 4
    for (int it = 0; it < SH_ITEMS; ++it) {
        const uint64_t i = tile_begin + (uint64_t)it * SH_THREADS + tid;
        uint32_t slot = SH_NONE;  // (segment, age) of the log entry this transaction becomes
        if (i < tile_end) {
            const uint32_t sv = __ldg(p.svc + i), jv = __ldg(p.jvm + i);
            const int32_t e = __ldg(p.el + i);
            const int64_t t = __ldg(p.ts + i);
            if (sv < p.num_services && jv < p.num_jvms && e >= 0 && t >= 0) {
                uint32_t sl = 0;
                long long latest;
                if (local) {
                    while (s_pos[sl] <= i) ++sl;
                    latest = s_lab[sl];
                } else {
                    sl = sh_ticks_le(p.ticks, seg_lo, seg_lo + nseg - 1, i) - seg_lo;
                    latest = seg_lo + sl == 0 ? p.latest0 : p.ticks[seg_lo + sl - 1].label;
                }
                const long long lab = apm_label(t, p.bucket_ms, p.inv_bucket);
                if (lab <= latest && lab >= latest - p.keep) {
                    slot = sl * APM_MAX_SLOTS + (uint32_t)(latest - lab);
                    my_max = (uint32_t)e > my_max ? (uint32_t)e : my_max;
                }
            }
        }
        // in-order streams put a whole warp on one counter: add once per warp then
        const uint32_t first = __shfl_sync(0xffffffffu, slot, 0);
        if (__all_sync(0xffffffffu, slot == first)) {
            if (lane == 0 && first != SH_NONE) {
                if (local) atomicAdd(&s_cnt[first], 32u);
                else atomicAdd(&p.counts[(size_t)(seg_lo + first / APM_MAX_SLOTS) * width + first % APM_MAX_SLOTS], 32u);
            }
        } else if (slot != SH_NONE) {
            if (local) atomicAdd(&s_cnt[slot], 1u);
            else atomicAdd(&p.counts[(size_t)(seg_lo + slot / APM_MAX_SLOTS) * width + slot % APM_MAX_SLOTS], 1u);
        }
    }
    my_max = __reduce_max_sync(0xffffffffu, my_max);
    if (lane == 0 && my_max) atomicMax(&s_max, my_max);
    __syncthreads();
    if (local)
        for (uint32_t j = tid; j < nseg * APM_MAX_SLOTS; j += SH_THREADS) {
            const uint32_t v = s_cnt[j], d = j % APM_MAX_SLOTS;
            if (v) atomicAdd(&p.counts[(size_t)(seg_lo + j / APM_MAX_SLOTS) * width + d], v);
        }
    if (tid == 0 && s_max) atomicMax(p.max_value, s_max);
}

// ---- K1: keyed segmented reduction ------------------------------------------------------------
// Each CTA keeps a hash table of (key, bucket-slot) partial accumulators in shared memory,
// streams its tiles of the segment through it with 128-bit coalesced loads, and flushes the
// table to the global accumulator planes with fire-and-forget L2 reductions.  Zipf-hot keys
// therefore cost native 32-bit shared-memory atomics instead of serialising on one L2 address;
// only table misses and durations >= 2^16 ms go to global memory directly.  All accumulators
// are integers, so the result does not depend on the order of the updates.
//
// Table entry (24 B, 4096 entries = 96 KB, two CTAs per SM), valid for values < 2^16 and at most
// 65535 updates between flushes:  tag | sum (u32, cannot overflow) | sq_lo (u32) |
// cnt:16 sq_hi:16 (carry of sq_lo) | min | max.
struct ReduceParams {
    const uint32_t* svc;
    const uint32_t* jvm;
    const int64_t* ts;
    const int32_t* el;
    uint64_t begin, end;
    int64_t bucket_ms;
    double inv_bucket;
    int64_t latest;
    int keep, S;
    int slot_latest;       // latest % S: slots are found from the bucket's age, not with a 64-bit modulo per transaction
    uint32_t num_services, num_jvms, K;
    AccSoA acc;
    uint8_t* seen;
    SlotMeta* slot_meta;
    uint64_t* const* logs;
    uint64_t log_cap;
    DevCounters* counters;
    int with_log;
    int vec_ok;            // all four column pointers are 16-byte aligned
    uint32_t tiles;        // tiles of K1_TILE transactions starting at (begin & ~3)
};

constexpr int K1_THREADS = 512;
constexpr int K1_PER_THREAD = 4;
constexpr int K1_TILE = K1_THREADS * K1_PER_THREAD;
constexpr int K1_TABLE = 4096;
constexpr int K1_PROBES = 4;
constexpr int K1_FLUSH_TILES = 65535 / K1_TILE;  // 31: keeps cnt < 2^16 and sum < 2^32 per entry
constexpr uint32_t K1_EMPTY = 0xFFFFFFFFu;
constexpr int K1_SLOT_SHIFT = 26;         // tag = slot << 26 | key  (K <= 2^26)

struct K1Smem {
    // used by both variants
    unsigned long long c[5];
    uint32_t vmax;
    uint32_t pad_;
    // the hash table: only the stats-only variant allocates it
    // one word longer than the table each: the fields of an entry then fall into different banks (see k1_chunk.cu)
    uint32_t tag[K1_TABLE + 1], sum[K1_TABLE + 1], sq_lo[K1_TABLE + 1], cnt_hi[K1_TABLE + 1], mn[K1_TABLE + 1], mx[K1_TABLE + 1];
};
constexpr size_t K1_SMEM_APPEND_ONLY = offsetof(K1Smem, tag);

__device__ __forceinline__ uint32_t k1_hash(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

__device__ __forceinline__ void k1_minmax_global(const AccSoA& acc, size_t ai, uint32_t lo, uint32_t hi) {
    // unconditional reductions: a read-then-update would stall on the load for no saving,
    // because most flushed entries are the first of their bucket
    atomicMin(&acc.vmin[ai], lo);
    atomicMax(&acc.vmax[ai], hi);
}

// one transaction straight to the global planes (table miss or a duration >= 2^16 ms)
__device__ __forceinline__ void k1_global_update(const ReduceParams& p, int slot, uint32_t key, uint32_t v) {
    const size_t ai = (size_t)slot * p.K + key;
    const unsigned long long sq = (unsigned long long)v * v;
    atomicAdd(&p.acc.cnt[ai], 1u);
    atomicAdd(&p.acc.sum[ai], (unsigned long long)v);
    atomicAdd(&p.acc.sq_lo[ai], sq & 0xFFFFFFFFull);
    if (sq >> 32) atomicAdd(&p.acc.sq_hi[ai], sq >> 32);
    k1_minmax_global(p.acc, ai, v, v);
}

__device__ __forceinline__ void k1_table_init(K1Smem& sm, int tid) {
    for (int i = tid; i < K1_TABLE; i += K1_THREADS) {
        sm.tag[i] = K1_EMPTY; sm.sum[i] = 0; sm.sq_lo[i] = 0; sm.cnt_hi[i] = 0; sm.mn[i] = 0xFFFFFFFFu; sm.mx[i] = 0;
    }
}

__device__ __forceinline__ void k1_table_flush(const ReduceParams& p, K1Smem& sm, int tid) {
    for (int i = tid; i < K1_TABLE; i += K1_THREADS) {
        const uint32_t tag = sm.tag[i];
        if (tag == K1_EMPTY) continue;
        const uint32_t k = tag & ((1u << K1_SLOT_SHIFT) - 1u);
        const size_t ai = (size_t)(tag >> K1_SLOT_SHIFT) * p.K + k;
        const uint32_t ch = sm.cnt_hi[i];
        atomicAdd(&p.acc.cnt[ai], ch >> 16);
        atomicAdd(&p.acc.sum[ai], (unsigned long long)sm.sum[i]);
        atomicAdd(&p.acc.sq_lo[ai], (unsigned long long)sm.sq_lo[i]);
        if (ch & 0xFFFFu) atomicAdd(&p.acc.sq_hi[ai], (unsigned long long)(ch & 0xFFFFu));
        k1_minmax_global(p.acc, ai, sm.mn[i], sm.mx[i]);
    }
}

template <bool TABLE>
__global__ void __launch_bounds__(K1_THREADS, 2) reduce_kernel_v3(const ReduceParams p) {
    extern __shared__ __align__(16) unsigned char k1_raw[];
    K1Smem& sm = *reinterpret_cast<K1Smem*>(k1_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    if (TABLE) k1_table_init(sm, tid);
    if (tid < 5) sm.c[tid] = 0;
    if (tid == 0) sm.vmax = 0;
    __syncthreads();

    const uint64_t i0 = p.begin & ~3ull;
    unsigned long long n_ing = 0, n_drop = 0, n_inv = 0, n_fut = 0, n_ovf = 0;
    uint32_t my_max = 0;
    int tiles_since_flush = 0;

    for (uint32_t tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        if (TABLE && tiles_since_flush == K1_FLUSH_TILES) {
            __syncthreads();
            k1_table_flush(p, sm, tid);
            __syncthreads();
            k1_table_init(sm, tid);
            __syncthreads();
            tiles_since_flush = 0;
        }
        ++tiles_since_flush;
        const uint64_t idx = i0 + (uint64_t)tile * K1_TILE + (uint64_t)tid * K1_PER_THREAD;
        uint32_t svc[4], jvm[4];
        int32_t el[4];
        int64_t ts[4];
        bool in[4];
        if (p.vec_ok && idx >= p.begin && idx + 4 <= p.end) {
            const uint4 s4 = __ldcs(reinterpret_cast<const uint4*>(p.svc + idx));
            const uint4 j4 = __ldcs(reinterpret_cast<const uint4*>(p.jvm + idx));
            const int4 e4 = __ldcs(reinterpret_cast<const int4*>(p.el + idx));
            const longlong2 t01 = __ldcs(reinterpret_cast<const longlong2*>(p.ts + idx));
            const longlong2 t23 = __ldcs(reinterpret_cast<const longlong2*>(p.ts + idx + 2));
            svc[0] = s4.x; svc[1] = s4.y; svc[2] = s4.z; svc[3] = s4.w;
            jvm[0] = j4.x; jvm[1] = j4.y; jvm[2] = j4.z; jvm[3] = j4.w;
            el[0] = e4.x; el[1] = e4.y; el[2] = e4.z; el[3] = e4.w;
            ts[0] = t01.x; ts[1] = t01.y; ts[2] = t23.x; ts[3] = t23.y;
            in[0] = in[1] = in[2] = in[3] = true;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint64_t i = idx + j;
                in[j] = i >= p.begin && i < p.end;
                svc[j] = in[j] ? __ldg(p.svc + i) : 0;
                jvm[j] = in[j] ? __ldg(p.jvm + i) : 0;
                el[j] = in[j] ? __ldg(p.el + i) : 0;
                ts[j] = in[j] ? __ldg(p.ts + i) : 0;
            }
        }
        int slot[4];       // -1: nothing to log
        uint32_t key[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            // lanes that took different paths through the last transaction (probe lengths, table misses) go on together:
            // left to themselves they stay split for the rest of the tile and the warp executes everything several times
            if (TABLE) __syncwarp();
            slot[j] = -1;
            key[j] = 0;
            if (!in[j]) continue;
            if (svc[j] >= p.num_services || jvm[j] >= p.num_jvms || el[j] < 0 || ts[j] < 0) { n_inv++; continue; }
            const int64_t lab = apm_label(ts[j], p.bucket_ms, p.inv_bucket);
            if (lab > p.latest) { n_fut++; continue; }
            const uint32_t k = jvm[j] * p.num_services + svc[j];
            // addData creates the (server, service) entry whatever the bucket's age (:125-126):
            // the key is emitted from now on even if this transaction is too old to count
            if (!p.seen[k]) p.seen[k] = 1;
            if (lab < p.latest - p.keep) { n_drop++; continue; }
            int sl = p.slot_latest - (int)(p.latest - lab);  // lab % S for latest - keep <= lab <= latest (keep < S)
            if (sl < 0) sl += p.S;
            const uint32_t v = (uint32_t)el[j];
            slot[j] = sl;
            key[j] = k;
            n_ing++;
            my_max = v > my_max ? v : my_max;
            if (!TABLE) continue;  // percentile mode: the accumulators are rebuilt from the sorted log
            bool done = false;
            if (v < 65536u) {
                // straight-line probing: a loop around the compare-and-swap keeps a split warp from ever rejoining (see k1_chunk.cu)
                const uint32_t tag = ((uint32_t)sl << K1_SLOT_SHIFT) | k;
                uint32_t h = k1_hash(tag) & (K1_TABLE - 1);
                int at = -1;
#pragma unroll
                for (int pr = 0; pr < K1_PROBES; ++pr) {
                    if (at < 0) {
                        uint32_t cur = sm.tag[h];
                        if (cur == K1_EMPTY) cur = atomicCAS(&sm.tag[h], K1_EMPTY, tag);
                        if (cur == tag || cur == K1_EMPTY) at = (int)h;
                        else h = (h + 1) & (K1_TABLE - 1);
                    }
                }
                if (at >= 0) {
                    atomicAdd(&sm.sum[at], v);
                    const uint32_t q = v * v;
                    const uint32_t o = atomicAdd(&sm.sq_lo[at], q);
                    atomicAdd(&sm.cnt_hi[at], 0x10000u + ((o + q < o) ? 1u : 0u));
                    if (v < sm.mn[at]) atomicMin(&sm.mn[at], v);
                    if (v > sm.mx[at]) atomicMax(&sm.mx[at], v);
                    done = true;
                }
            }
            if (!done) k1_global_update(p, sl, k, v);
        }

        if (!TABLE) {
            // Append (key, value) words to their bucket's log.  Every warp reserves its own stretch of the log - one
            // atomicAdd per (warp, slot), normally one per 128 transactions - and writes it as 1 KB of consecutive
            // words, so the tile loop has no block barrier and no warp waits for another one's reservation.
            uint32_t pend = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) pend |= slot[j] >= 0 ? (1u << j) : 0u;
            for (;;) {
                uint32_t mine = 0xFFFFFFFFu;
#pragma unroll
                for (int j = 3; j >= 0; --j)
                    if (pend & (1u << j)) mine = ((uint32_t)(lane * 4 + j) << 8) | (uint32_t)slot[j];
                const uint32_t wmin = __reduce_min_sync(0xffffffffu, mine);
                if (wmin == 0xFFFFFFFFu) break;
                const int gslot = (int)(wmin & 0xFFu);
                uint32_t c = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) c += ((pend >> j) & 1u) && slot[j] == gslot ? 1u : 0u;
                uint32_t incl = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += y;
                }
                const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(&p.slot_meta[gslot].count, (unsigned long long)total);
                base = __shfl_sync(0xffffffffu, base, 0);
                unsigned long long pos = base + incl - c;
                uint64_t* lg = p.logs[gslot];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (((pend >> j) & 1u) && slot[j] == gslot) {
                        if (pos < p.log_cap) lg[pos] = ((uint64_t)key[j] << 32) | (uint32_t)el[j];
                        else n_ovf++;
                        ++pos;
                        pend &= ~(1u << j);
                    }
                }
            }
        }
    }

    __syncthreads();
    if (TABLE) k1_table_flush(p, sm, tid);
    my_max = __reduce_max_sync(0xffffffffu, my_max);
    if (lane == 0 && my_max) atomicMax(&sm.vmax, my_max);
    {
        // per-warp totals first: a 64-bit shared atomicAdd is a compare-and-swap loop on sm_100a, and 512 threads spinning on
        // one word were 45 % of this kernel's stall samples
        const unsigned long long mine[5] = {n_ing, n_drop, n_inv, n_fut, n_ovf};
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            const uint32_t wsum = __reduce_add_sync(0xffffffffu, (uint32_t)mine[q]);  // a thread sees < 2^32 transactions
            if (lane == 0 && wsum) atomicAdd(&sm.c[q], (unsigned long long)wsum);
        }
    }
    __syncthreads();
    if (tid == 0) {
        if (sm.c[0]) atomicAdd(&p.counters->ingested, sm.c[0]);
        if (sm.c[1]) atomicAdd(&p.counters->dropped, sm.c[1]);
        if (sm.c[2]) atomicAdd(&p.counters->invalid, sm.c[2]);
        if (sm.c[3]) atomicAdd(&p.counters->future, sm.c[3]);
        if (sm.c[4]) atomicAdd(&p.counters->log_overflow, sm.c[4]);
        if (sm.vmax > p.counters->max_value) atomicMax(&p.counters->max_value, sm.vmax);
    }
}

}  // namespace

// ---- the pre-pass of a chunk: where its ticks are and what every segment appends to every bucket ------------
// begin:  tick finder on `st`, count + first slice of the marks on their way to the host, pp.ev recorded behind them.
// finish: waits for pp.ev, sorts the marks by position, runs the segment table kernel and brings it back.
// Between the two the host is free to queue other work (the ticks of the previous chunk).
constexpr unsigned int PP_FIRST_MARKS = 256;

apm_status apm_prepass_begin(apm_ctx* c, apm_ctx::PrePass& pp, const apm_tx_columns& d, uint64_t n, int64_t latest0, cudaStream_t st) {
    ProfScope prof_scope_(c, APM_PG_TICKFIND, st);
    pp.nt = 0;
    pp.seg_max = 0;
    pp.latest0 = latest0;
    pp.external = false;
    *pp.h_nticks = 0;
    if (c->ext_armed) {
        // The host has located the ticks: this chunk takes the marks that fall inside it (the call's last chunk also the
        // ones behind its last transaction), rebased to the chunk.  No tick finder runs.
        pp.external = true;
        unsigned int k = 0;
        const uint64_t lo = c->ext_off, hi = c->ext_off + n;
        int64_t prev = latest0;
        while (c->ext_next < c->ext_marks.size()) {
            const TickMark& m = c->ext_marks[c->ext_next];
            if (m.pos > hi || (m.pos == hi && !c->ext_last)) break;
            if (m.pos < lo || m.label <= prev)
                return apm_fail(c, APM_E_INVALID, "tick marks must be ascending in position and strictly ascending in label, above latestBucket");
            if (k >= c->max_ticks_per_chunk) return apm_fail(c, APM_E_CAPACITY, "too many ticks in one chunk");
            pp.h_ticks[k].pos = m.pos - lo;
            pp.h_ticks[k].label = m.label;
            prev = m.label;
            ++k;
            ++c->ext_next;
        }
        *pp.h_nticks = k;
        CUDA_TRY(c, cudaEventRecord(pp.ev, st));
        return APM_OK;
    }
    if (n == 0) { CUDA_TRY(c, cudaEventRecord(pp.ev, st)); return APM_OK; }
    const uint64_t ntiles = (n + TF_TILE - 1) / TF_TILE;
    if (ntiles > c->tf_tiles_cap) return apm_fail(c, APM_E_CAPACITY, "chunk larger than max_batch");
    CUDA_TRY(c, cudaMemsetAsync(pp.d_nticks, 0, sizeof(unsigned int), st));
    tf_tile_max_kernel<<<(unsigned)ntiles, TF_THREADS, 0, st>>>(d.end_ts_ms, n, c->cfg.bucket_ms, c->inv_bucket, pp.tf_tile_max);
    tf_scan_kernel<<<1, 1024, 0, st>>>(pp.tf_tile_max, ntiles, latest0, pp.tf_tile_prefix);
    tf_emit_kernel<<<(unsigned)ntiles, TF_THREADS, 0, st>>>(d.end_ts_ms, n, c->cfg.bucket_ms, c->inv_bucket, pp.tf_tile_max,
                                                           pp.tf_tile_prefix, pp.d_ticks, pp.d_nticks, c->max_ticks_per_chunk);
    c->launches += 3;
    CUDA_TRY(c, cudaGetLastError());
    // the count and a first slice of the marks come back in one round trip; only a chunk with more ticks than
    // the slice holds pays a second one
    const unsigned int first = c->max_ticks_per_chunk < PP_FIRST_MARKS ? c->max_ticks_per_chunk : PP_FIRST_MARKS;
    CUDA_TRY(c, cudaMemcpyAsync(pp.h_nticks, pp.d_nticks, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(c, cudaMemcpyAsync(pp.h_ticks, pp.d_ticks, first * sizeof(TickMark), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(c, cudaEventRecord(pp.ev, st));
    return APM_OK;
}

// the segment table of a chunk whose (sorted) marks are in pp.h_ticks[0 .. pp.nt): pp.h_seg_counts, pp.seg_max
static apm_status seg_table(apm_ctx* c, apm_ctx::PrePass& pp, const apm_tx_columns& d, uint64_t n, cudaStream_t st) {
    ProfScope prof_scope_(c, APM_PG_TICKFIND, st);
    const unsigned int nt = pp.nt;
    const size_t cells = (size_t)(nt + 1) * (c->keep + 1);
    memset(pp.h_seg_counts, 0, cells * sizeof(uint32_t));
    pp.seg_max = 0;
    if (n == 0) return APM_OK;
    if (nt) CUDA_TRY(c, cudaMemcpyAsync(pp.d_ticks, pp.h_ticks, nt * sizeof(TickMark), cudaMemcpyHostToDevice, st));
    CUDA_TRY(c, cudaMemsetAsync(pp.d_seg_counts, 0, (cells + 1) * sizeof(uint32_t), st));
    SegHistParams p;
    p.svc = d.service_id; p.jvm = d.jvm_id; p.ts = d.end_ts_ms; p.el = d.elapsed_ms; p.n = n;
    p.bucket_ms = c->cfg.bucket_ms; p.inv_bucket = c->inv_bucket; p.latest0 = pp.latest0; p.keep = c->keep;
    p.num_services = c->cfg.num_services; p.num_jvms = c->cfg.num_jvms;
    p.ticks = pp.d_ticks; p.nt = nt;
    p.counts = pp.d_seg_counts; p.max_value = pp.d_seg_counts + cells;  // the chunk maximum rides behind the table
    seg_hist_kernel<<<(unsigned)((n + SH_TILE - 1) / SH_TILE), SH_THREADS, 0, st>>>(p);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    CUDA_TRY(c, cudaMemcpyAsync(pp.h_seg_counts, pp.d_seg_counts, (cells + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(c, cudaStreamSynchronize(st));
    pp.seg_max = pp.h_seg_counts[cells];
    c->counters.d2h_bytes += (cells + 1) * sizeof(uint32_t);
    c->counters.h2d_bytes += nt * sizeof(TickMark);
    return APM_OK;
}

apm_status apm_prepass_finish(apm_ctx* c, apm_ctx::PrePass& pp, const apm_tx_columns& d, uint64_t n, cudaStream_t st) {
    CUDA_TRY(c, cudaEventSynchronize(pp.ev));
    const unsigned int nt = *pp.h_nticks;
    if (nt > c->max_ticks_per_chunk) return apm_fail(c, APM_E_CAPACITY, "too many ticks in one chunk");
    pp.nt = nt;
    if (!pp.external) {
        const unsigned int first = c->max_ticks_per_chunk < PP_FIRST_MARKS ? c->max_ticks_per_chunk : PP_FIRST_MARKS;
        if (nt > first) {
            CUDA_TRY(c, cudaMemcpyAsync(pp.h_ticks + first, pp.d_ticks + first, (nt - first) * sizeof(TickMark), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(c, cudaStreamSynchronize(st));
        }
        c->counters.d2h_bytes += sizeof(unsigned int) + (nt > first ? nt : first) * sizeof(TickMark);
        std::sort(pp.h_ticks, pp.h_ticks + nt, [](const TickMark& a, const TickMark& b) { return a.pos < b.pos; });
    }
    if ((c->cfg.flags & APM_F_PERCENTILES) && c->cfg.num_percentiles > 0) return seg_table(c, pp, d, n, st);
    if (nt && c->stats_chunk) {
        // the whole-chunk stats path (k1_chunk.cu) reads the marks on the device, in position order
        CUDA_TRY(c, cudaMemcpyAsync(pp.d_ticks, pp.h_ticks, nt * sizeof(TickMark), cudaMemcpyHostToDevice, st));
        CUDA_TRY(c, cudaStreamSynchronize(st));
        c->counters.h2d_bytes += nt * sizeof(TickMark);
    }
    return APM_OK;
}

// a batch that is known to hold no tick (apm_ingest_commit): one segment under latest0
apm_status apm_prepass_segment(apm_ctx* c, apm_ctx::PrePass& pp, const apm_tx_columns& d, uint64_t n, int64_t latest0, cudaStream_t st) {
    pp.nt = 0;
    pp.latest0 = latest0;
    return seg_table(c, pp, d, n, st);
}

apm_status apm_launch_reduce(apm_ctx* c, const apm_tx_columns& d, uint64_t begin, uint64_t end) {
    ProfScope prof_scope_(c, APM_PG_REDUCE);
    if (end <= begin) return APM_OK;
    ReduceParams p;
    p.svc = d.service_id; p.jvm = d.jvm_id; p.ts = d.end_ts_ms; p.el = d.elapsed_ms;
    p.begin = begin; p.end = end;
    p.bucket_ms = c->cfg.bucket_ms; p.inv_bucket = c->inv_bucket; p.latest = c->latest;
    p.keep = c->keep; p.S = c->S;
    p.slot_latest = (int)(c->latest % c->S);
    p.num_services = c->cfg.num_services; p.num_jvms = c->cfg.num_jvms; p.K = c->K;
    p.acc = c->acc; p.seen = c->seen; p.slot_meta = c->slot_meta; p.logs = c->d_log_ptrs; p.log_cap = c->log_cap;
    p.counters = c->d_counters;
    p.with_log = (c->cfg.flags & APM_F_PERCENTILES) && c->cfg.num_percentiles > 0;
    const uintptr_t align = (uintptr_t)d.service_id | (uintptr_t)d.jvm_id | (uintptr_t)d.end_ts_ms | (uintptr_t)d.elapsed_ms;
    p.vec_ok = (align & 15) == 0;
    const uint64_t i0 = begin & ~3ull;
    const uint64_t tiles = (end - i0 + K1_TILE - 1) / K1_TILE;
    if (tiles > 0xFFFFFFFFull) return apm_fail(c, APM_E_CAPACITY, "segment too large");
    p.tiles = (uint32_t)tiles;
    // every CTA flushes its table once: give each at least k1_tiles_per_cta tiles, cap at 2 CTAs per SM
    uint64_t blocks = (tiles + c->k1_tiles_per_cta - 1) / c->k1_tiles_per_cta;
    const uint64_t max_blocks = (uint64_t)c->num_sms * 2;
    if (blocks > max_blocks) blocks = max_blocks;
    if (blocks < 1) blocks = 1;
    static unsigned long long attr_done = 0;  // per device: the limit belongs to the device that is current when it is set
    if (apm_first_on_device(&attr_done)) {
        CUDA_TRY(c, cudaFuncSetAttribute(reduce_kernel_v3<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K1Smem)));
        CUDA_TRY(c, cudaFuncSetAttribute(reduce_kernel_v3<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K1Smem)));
    }
    if (p.with_log) {
        // append-only: no table in shared memory and none to flush, so the tiles spread over every SM (2 CTAs each: 64 registers per thread)
        const uint64_t cap4 = (uint64_t)c->num_sms * 2;
        uint64_t b2 = tiles < cap4 ? tiles : cap4;
        reduce_kernel_v3<false><<<(unsigned)b2, K1_THREADS, K1_SMEM_APPEND_ONLY, c->s_main>>>(p);
    } else {
        reduce_kernel_v3<true><<<(unsigned)blocks, K1_THREADS, sizeof(K1Smem), c->s_main>>>(p);
    }
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k1_ingest() {
    const void* fns[] = {(const void*)reduce_kernel_v3<true>, (const void*)reduce_kernel_v3<false>, (const void*)seg_hist_kernel, (const void*)tf_emit_kernel, (const void*)tf_scan_kernel, (const void*)tf_tile_max_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
