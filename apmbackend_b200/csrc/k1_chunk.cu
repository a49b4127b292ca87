// K1c: the stats-only path (no percentiles, no z-score: BASELINE config 2) a whole chunk at a time.
//
// consumeMsg (stream_calc_stats.js:331-371) alternates addData and ticks, and the per-segment path mirrors that: one
// keyed-reduce launch, one slot clear and one window kernel per tick, and - what made it 6 % of the HBM roofline - a
// flush of every CTA's shared-memory table at EVERY tick, because a launch covered one segment.  Nothing in this mode
// needs the device to stop at a tick: all accumulators are order-free integers and keys are independent.  So:
//
//   reduce   ONE launch per chunk.  Every CTA streams a contiguous stretch of the chunk through its shared-memory table
//            of (bucket age, key) partials and flushes it once, into accumulator planes of ITS segment - the chunk's
//            tick marks are on the device already, a transaction's segment is a look at the next mark.  Planes are
//            [segment][age][key], so what each tick may see stays separable.
//   ticks    ONE launch per chunk, one thread per key: for tick i it folds segment i's planes into the key's bucket
//            accumulators, clears the slots the tick evicts (removeOldBuckets, :103-113), combines the window and
//            writes the stat record (generateAllStatsToQueue, :157-203) - the same order as the per-segment path, so the
//            same records bit for bit - then goes on to tick i + 1.
#include <vector>

#include "ctx.h"

namespace {

constexpr int KC_THREADS = 1024;
constexpr int KC_PER_THREAD = 4;
constexpr int KC_TILE = KC_THREADS * KC_PER_THREAD;  // 4096 transactions
constexpr uint32_t KC_TABLE = 7168;                  // entries of 28 B: 196 KB, one CTA per SM
constexpr int KC_PROBES = 4;
constexpr uint32_t KC_EMPTY = 0xFFFFFFFFu;
constexpr int KC_AGE_SHIFT = 26;                     // tag = age << 26 | key (K <= 2^26, age < 64)
constexpr uint32_t KC_NO_SEG = 0xFFFFFFFFu;

struct SegPlanes {  // each [segments][keep + 1][K]
    unsigned long long* sum;
    unsigned long long* sq_lo;
    unsigned long long* sq_hi;
    uint32_t* cnt;
    uint32_t* vmin;
    uint32_t* vmax;
};

// Table entry without update limits: a 32-bit sum and a 32-bit sum of squares with their carries counted beside them
// (values below 2^16, so a square fits 32 bits and one update carries at most once).
// Each field array is one word longer than the table: field f of entry h then sits in bank (h + f) mod 32.  With arrays
// of exactly KC_TABLE words (a multiple of 32) all seven words of an entry share ONE bank, and a Zipf-hot key's five
// atomics per transaction queue behind each other there.
struct KcHead {
    unsigned long long c[4];   // ingested, dropped, invalid, future
    uint32_t cur_seg, next_pos_lo, next_pos_hi, pad_;
    long long cur_latest;
};
struct KcSmem {
    KcHead h;
    uint32_t tag[KC_TABLE + 1], cnt[KC_TABLE + 1], sum_lo[KC_TABLE + 1], sq_lo[KC_TABLE + 1], carry[KC_TABLE + 1] /* sq: bits 8.., sum: bits 0..7 */,
        mn[KC_TABLE + 1], mx[KC_TABLE + 1];
};

// When every key has room in shared memory (K <= KC_DIRECT_KEYS: the reference's 400 keys, BASELINE config 2's 10 000),
// the table needs no tags, no hash and no probing: entry = key, for the segment's latest bucket (age 0; the few late
// arrivals go straight to the planes).  20 bytes per key: count, sum, sum of squares (32 bits each, carries counted in a
// fourth word) and two 16-bit FILTERS for min / max - a duration below the filter (above, for max) goes to the planes'
// own atomicMin / atomicMax and is stored into the filter with a plain 16-bit store.  A filter is never below the minimum
// the planes hold (only values that were also sent are stored), so nothing that matters is filtered out; a lost race
// only lets a few more values through.
constexpr uint32_t KC_DIRECT_KEYS = 11584;
struct KcDirect {
    KcHead h;
    uint32_t words[1];  // cnt[Kp] sum_lo[Kp] sq_lo[Kp] carry[Kp], then uint16 mn[Kp] mx[Kp];  Kp = K rounded up to 32
};
__host__ __device__ inline uint32_t kc_direct_kp(uint32_t K) { return (K + 31u) & ~31u; }
__host__ __device__ inline size_t kc_direct_bytes(uint32_t K) { return sizeof(KcHead) + (size_t)kc_direct_kp(K) * 20u; }

struct ChunkReduceParams {
    const uint32_t* svc;
    const uint32_t* jvm;
    const int64_t* ts;
    const int32_t* el;
    uint64_t n;
    int64_t bucket_ms;
    double inv_bucket;
    int64_t latest0;
    int keep;
    uint32_t num_services, num_jvms, K;
    const TickMark* marks;  // sorted by position
    uint32_t nt;
    SegPlanes planes;
    uint32_t* first_seg;    // [K] first segment of this chunk in which the key was named (addData creates the entry, :125-126)
    DevCounters* counters;
    uint32_t tiles;
    int vec_ok;             // all four column pointers are 16-byte aligned
};

__device__ __forceinline__ uint32_t kc_hash(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

__device__ __forceinline__ uint32_t kc_marks_le(const TickMark* m, uint32_t lo, uint32_t hi, uint64_t i) {
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (m[mid].pos <= i) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ size_t kc_cell(const ChunkReduceParams& p, uint32_t seg, uint32_t age, uint32_t key) {
    return ((size_t)seg * (size_t)(p.keep + 1) + age) * p.K + key;
}

__device__ __forceinline__ void kc_direct(const ChunkReduceParams& p, uint32_t seg, uint32_t age, uint32_t key, uint32_t v) {
    const size_t ai = kc_cell(p, seg, age, key);
    const unsigned long long sq = (unsigned long long)v * v;
    atomicAdd(&p.planes.cnt[ai], 1u);
    atomicAdd(&p.planes.sum[ai], (unsigned long long)v);
    atomicAdd(&p.planes.sq_lo[ai], sq & 0xFFFFFFFFull);
    if (sq >> 32) atomicAdd(&p.planes.sq_hi[ai], sq >> 32);
    atomicMin(&p.planes.vmin[ai], v);
    atomicMax(&p.planes.vmax[ai], v);
}

__device__ __forceinline__ void kc_table_init(KcSmem& sm, int tid) {
    for (uint32_t i = tid; i < KC_TABLE; i += KC_THREADS) {
        sm.tag[i] = KC_EMPTY; sm.cnt[i] = 0; sm.sum_lo[i] = 0; sm.sq_lo[i] = 0; sm.carry[i] = 0; sm.mn[i] = 0xFFFFFFFFu; sm.mx[i] = 0;
    }
}

__device__ __forceinline__ void kc_table_flush(const ChunkReduceParams& p, KcSmem& sm, int tid, uint32_t seg) {
    for (uint32_t i = tid; i < KC_TABLE; i += KC_THREADS) {
        const uint32_t tag = sm.tag[i];
        if (tag == KC_EMPTY) continue;
        const size_t ai = kc_cell(p, seg, tag >> KC_AGE_SHIFT, tag & ((1u << KC_AGE_SHIFT) - 1u));
        const uint32_t cy = sm.carry[i];
        atomicAdd(&p.planes.cnt[ai], sm.cnt[i]);
        atomicAdd(&p.planes.sum[ai], (unsigned long long)sm.sum_lo[i] + ((unsigned long long)(cy & 0xFFu) << 32));
        atomicAdd(&p.planes.sq_lo[ai], (unsigned long long)sm.sq_lo[i]);
        if (cy >> 8) atomicAdd(&p.planes.sq_hi[ai], (unsigned long long)(cy >> 8));
        atomicMin(&p.planes.vmin[ai], sm.mn[i]);
        atomicMax(&p.planes.vmax[ai], sm.mx[i]);
    }
}

template <bool DIRECT>
__global__ void __launch_bounds__(KC_THREADS, 1) reduce_chunk_kernel(const ChunkReduceParams p) {
    extern __shared__ __align__(16) unsigned char kc_raw[];
    KcSmem& sm = *reinterpret_cast<KcSmem*>(kc_raw);
    KcHead& hd = *reinterpret_cast<KcHead*>(kc_raw);  // first member of either layout
    const int tid = threadIdx.x, lane = tid & 31;
    // direct layout (DIRECT only)
    const uint32_t Kp = kc_direct_kp(p.K);
    uint32_t* const d_cnt = reinterpret_cast<KcDirect*>(kc_raw)->words;
    uint32_t* const d_sum = d_cnt + Kp;
    uint32_t* const d_sq = d_sum + Kp;
    uint32_t* const d_carry = d_sq + Kp;
    unsigned short* const d_mn = reinterpret_cast<unsigned short*>(d_carry + Kp);
    unsigned short* const d_mx = d_mn + Kp;
    auto table_init = [&]() {
        if constexpr (DIRECT) {
            for (uint32_t i = tid; i < Kp; i += KC_THREADS) { d_cnt[i] = 0; d_sum[i] = 0; d_sq[i] = 0; d_carry[i] = 0; d_mn[i] = 0xFFFFu; d_mx[i] = 0; }
        } else {
            kc_table_init(sm, tid);
        }
    };
    auto table_flush = [&](uint32_t seg) {
        if constexpr (DIRECT) {
            for (uint32_t k = tid; k < p.K; k += KC_THREADS) {
                const uint32_t c = d_cnt[k];
                if (!c) continue;
                const size_t ai = kc_cell(p, seg, 0, k);  // min / max are in the planes already
                const uint32_t cy = d_carry[k];
                atomicAdd(&p.planes.cnt[ai], c);
                atomicAdd(&p.planes.sum[ai], (unsigned long long)d_sum[k] + ((unsigned long long)(cy & 0xFFu) << 32));
                atomicAdd(&p.planes.sq_lo[ai], (unsigned long long)d_sq[k]);
                if (cy >> 8) atomicAdd(&p.planes.sq_hi[ai], (unsigned long long)(cy >> 8));
            }
        } else {
            kc_table_flush(p, sm, tid, seg);
        }
    };
    table_init();
    if (tid < 4) hd.c[tid] = 0;
    // this CTA's contiguous stretch of tiles
    const uint32_t t0 = (uint32_t)(((uint64_t)blockIdx.x * p.tiles) / gridDim.x);
    const uint32_t t1 = (uint32_t)(((uint64_t)(blockIdx.x + 1) * p.tiles) / gridDim.x);
    if (tid == 0) {
        const uint32_t s = kc_marks_le(p.marks, 0, p.nt, (uint64_t)t0 * KC_TILE);
        const unsigned long long np = s < p.nt ? p.marks[s].pos : ~0ull;
        hd.cur_seg = s;
        hd.next_pos_lo = (uint32_t)np; hd.next_pos_hi = (uint32_t)(np >> 32);
        hd.cur_latest = s == 0 ? p.latest0 : p.marks[s - 1].label;
    }
    __syncthreads();
    unsigned long long n_ing = 0, n_drop = 0, n_inv = 0, n_fut = 0;
    uint32_t since_flush = 0;  // a CTA-uniform count of tiles: the sum carry has 8 bits (2^40 / 2^16 values = 2^24 updates per entry)
    for (uint32_t tile = t0; tile < t1; ++tile) {
        const uint64_t tile_begin = (uint64_t)tile * KC_TILE;
        unsigned long long next_pos = ((unsigned long long)hd.next_pos_hi << 32) | hd.next_pos_lo;
        if (tile_begin >= next_pos || since_flush >= 4000u) {
            // the stretch enters a new segment (or the carries near their width): hand the table to the planes of the old one
            __syncthreads();
            table_flush(hd.cur_seg);
            __syncthreads();
            table_init();
            if (tid == 0) {
                const uint32_t s = kc_marks_le(p.marks, hd.cur_seg, p.nt, tile_begin);
                const unsigned long long np = s < p.nt ? p.marks[s].pos : ~0ull;
                hd.cur_seg = s;
                hd.next_pos_lo = (uint32_t)np; hd.next_pos_hi = (uint32_t)(np >> 32);
                hd.cur_latest = s == 0 ? p.latest0 : p.marks[s - 1].label;
            }
            __syncthreads();
            next_pos = ((unsigned long long)hd.next_pos_hi << 32) | hd.next_pos_lo;
            since_flush = 0;
        }
        ++since_flush;
        const uint32_t cur_seg = hd.cur_seg;
        const long long cur_latest = hd.cur_latest;
        const uint64_t idx = tile_begin + (uint64_t)tid * KC_PER_THREAD;
        uint32_t svc[4], jvm[4];
        int32_t el[4];
        int64_t ts[4];
        bool in[4];
        if (p.vec_ok && idx + 4 <= p.n) {
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
                in[j] = i < p.n;
                svc[j] = in[j] ? p.svc[i] : 0; jvm[j] = in[j] ? p.jvm[i] : 0; el[j] = in[j] ? p.el[i] : 0; ts[j] = in[j] ? p.ts[i] : 0;
            }
        }
        // DIRECT: in a tile without a tick, a transaction of the segment's latest bucket with a 16-bit duration - nearly
        // all of them - is four compares away from its table entry (3 % on cfg2: the kernel waits for its loads more than
        // it computes; 512 threads with the next tile's loads in flight were slower, 16 warps do not fill the SM)
        const bool plain_tile = tile_begin + KC_TILE <= next_pos;
        const int64_t b0_tile = cur_latest * p.bucket_ms;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (DIRECT && in[j] && plain_tile && svc[j] < p.num_services && jvm[j] < p.num_jvms && (uint32_t)el[j] < 65536u &&
                (uint64_t)(ts[j] - b0_tile) < (uint64_t)p.bucket_ms) {
                const uint32_t k = jvm[j] * p.num_services + svc[j], v = (uint32_t)el[j];
                n_ing++;
                const uint32_t oc = atomicAdd(&d_cnt[k], 1u);
                const uint32_t os = atomicAdd(&d_sum[k], v);
                const uint32_t q = v * v;
                const uint32_t oq = atomicAdd(&d_sq[k], q);
                const uint32_t cy = ((os + v < os) ? 1u : 0u) | ((oq + q < oq) ? 0x100u : 0u);
                if (cy) atomicAdd(&d_carry[k], cy);
                if (oc == 0 || v < d_mn[k] || v > d_mx[k]) {
                    const size_t ai = kc_cell(p, cur_seg, 0, k);
                    if (v < d_mn[k]) { atomicMin(&p.planes.vmin[ai], v); d_mn[k] = (unsigned short)v; }
                    if (v > d_mx[k]) { atomicMax(&p.planes.vmax[ai], v); d_mx[k] = (unsigned short)v; }
                    if (oc == 0 && p.first_seg[k] > cur_seg) atomicMin(&p.first_seg[k], cur_seg);
                }
            } else {
            // what becomes of transaction j: nothing (skipped / counted only), a direct update of the planes, or a table update
            uint32_t tag = KC_EMPTY, v = 0, seg = cur_seg, age = 0, k = 0;
            bool direct = false;
            if (in[j]) {
                if (svc[j] >= p.num_services || jvm[j] >= p.num_jvms || el[j] < 0 || ts[j] < 0) {
                    n_inv++;
                } else {
                    // the segment of this transaction: the CTA's current one unless the tile straddles a tick
                    long long latest = cur_latest;
                    if (idx + j >= next_pos) {
                        seg = kc_marks_le(p.marks, cur_seg, p.nt, idx + j);
                        latest = p.marks[seg - 1].label;
                    }
                    // most transactions fall into the segment's latest bucket: two compares instead of the division
                    const int64_t b0 = latest * p.bucket_ms;
                    const int64_t lab = (ts[j] >= b0 && ts[j] - b0 < p.bucket_ms) ? (int64_t)latest : apm_label(ts[j], p.bucket_ms, p.inv_bucket);
                    if (lab > latest) {
                        n_fut++;
                    } else {
                        k = jvm[j] * p.num_services + svc[j];
                        if (lab < latest - p.keep) {
                            n_drop++;
                            if (p.first_seg[k] > seg) atomicMin(&p.first_seg[k], seg);  // too old to count, but it names the key (:125-126)
                        } else {
                            age = (uint32_t)(latest - lab);
                            v = (uint32_t)el[j];
                            n_ing++;
                            if (seg == cur_seg && v < 65536u && (!DIRECT || age == 0)) tag = (age << KC_AGE_SHIFT) | k;
                            else direct = true;
                        }
                    }
                }
            }
            bool done = tag == KC_EMPTY && !direct;
            if (DIRECT && tag != KC_EMPTY) {
                const uint32_t oc = atomicAdd(&d_cnt[k], 1u);
                const uint32_t os = atomicAdd(&d_sum[k], v);
                const uint32_t q = v * v;
                const uint32_t oq = atomicAdd(&d_sq[k], q);
                const uint32_t cy = ((os + v < os) ? 1u : 0u) | ((oq + q < oq) ? 0x100u : 0u);
                if (cy) atomicAdd(&d_carry[k], cy);
                if (oc == 0 || v < d_mn[k] || v > d_mx[k]) {
                    const size_t ai = kc_cell(p, seg, 0, k);
                    if (v < d_mn[k]) { atomicMin(&p.planes.vmin[ai], v); d_mn[k] = (unsigned short)v; }
                    if (v > d_mx[k]) { atomicMax(&p.planes.vmax[ai], v); d_mx[k] = (unsigned short)v; }
                    // whoever touches the entry first names the key for this segment: once per (CTA, segment, key)
                    if (oc == 0 && p.first_seg[k] > seg) atomicMin(&p.first_seg[k], seg);
                }
                done = true;
            } else if (tag != KC_EMPTY) {
                // Straight-line probing (no loop around the compare-and-swap): with a backward branch here the compiler
                // treats the lanes as possibly waiting for each other and never lets a split warp rejoin - every later
                // transaction of the tile was then executed once per group of lanes.
                uint32_t h = (uint32_t)(((unsigned long long)kc_hash(tag) * KC_TABLE) >> 32);
                int slot = -1;
#pragma unroll
                for (int pr = 0; pr < KC_PROBES; ++pr) {
                    if (slot < 0) {
                        uint32_t cur = sm.tag[h];
                        if (cur == KC_EMPTY) {
                            cur = atomicCAS(&sm.tag[h], KC_EMPTY, tag);
                            // whoever creates the entry names the key for this segment: once per (CTA, segment, key), not per transaction
                            if (cur == KC_EMPTY && p.first_seg[k] > seg) atomicMin(&p.first_seg[k], seg);
                        }
                        if (cur == tag || cur == KC_EMPTY) slot = (int)h;
                        else h = h + 1 == KC_TABLE ? 0 : h + 1;
                    }
                }
                if (slot >= 0) {
                    atomicAdd(&sm.cnt[slot], 1u);
                    const uint32_t os = atomicAdd(&sm.sum_lo[slot], v);
                    const uint32_t q = v * v;
                    const uint32_t oq = atomicAdd(&sm.sq_lo[slot], q);
                    const uint32_t cy = ((os + v < os) ? 1u : 0u) | ((oq + q < oq) ? 0x100u : 0u);
                    if (cy) atomicAdd(&sm.carry[slot], cy);
                    if (v < sm.mn[slot]) atomicMin(&sm.mn[slot], v);
                    if (v > sm.mx[slot]) atomicMax(&sm.mx[slot], v);
                    done = true;
                }
            }
            if (!done) {
                kc_direct(p, seg, age, k, v);
                if (p.first_seg[k] > seg) atomicMin(&p.first_seg[k], seg);
            }
            }
            __syncwarp();  // lanes that took different paths (probe lengths, misses) go on together
        }
    }
    __syncthreads();
    table_flush(hd.cur_seg);
    const unsigned long long mine[4] = {n_ing, n_drop, n_inv, n_fut};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint32_t wsum = __reduce_add_sync(0xffffffffu, (uint32_t)mine[q]);
        if (lane == 0 && wsum) atomicAdd(&hd.c[q], (unsigned long long)wsum);
    }
    __syncthreads();
    if (tid == 0) {
        if (hd.c[0]) atomicAdd(&p.counters->ingested, hd.c[0]);
        if (hd.c[1]) atomicAdd(&p.counters->dropped, hd.c[1]);
        if (hd.c[2]) atomicAdd(&p.counters->invalid, hd.c[2]);
        if (hd.c[3]) atomicAdd(&p.counters->future, hd.c[3]);
    }
}

// ---- the chunk's ticks, one thread per key ------------------------------------------------------------------------
struct TickDesc {
    long long tick_ts;          // edgeTs of the tick
    unsigned long long clear;   // slots the tick evicts / re-labels
    int slot_latest_before;     // (latestBucket in force during the segment before the tick) mod S
    int nwin;
    unsigned char win[APM_MAX_SLOTS];  // window slots, ascending label
};

struct ChunkTickParams {
    AccSoA acc;                 // [S][K] bucket accumulators
    SegPlanes planes;
    const uint32_t* first_seg;
    uint8_t* seen;
    const TickDesc* desc;       // [nt]
    uint32_t nt, K;
    int keep, S;
    int slot_latest_final;      // latestBucket mod S after the last tick: the tail segment goes there
    double tpm_div;
    apm_stat_record* stats;     // [nt][K] or nullptr
    uint32_t* unseen_records;
};

__device__ __forceinline__ void kc_fold(const ChunkTickParams& p, uint32_t seg, int slot_latest, uint32_t key) {
    for (int age = 0; age <= p.keep; ++age) {
        const size_t ci = ((size_t)seg * (size_t)(p.keep + 1) + age) * p.K + key;
        const uint32_t c = p.planes.cnt[ci];
        if (!c) continue;
        int sl = slot_latest - age;
        if (sl < 0) sl += p.S;
        const size_t ai = (size_t)sl * p.K + key;
        p.acc.cnt[ai] += c;
        p.acc.sum[ai] += p.planes.sum[ci];
        p.acc.sq_lo[ai] += p.planes.sq_lo[ci];
        p.acc.sq_hi[ai] += p.planes.sq_hi[ci];
        const uint32_t mn = p.planes.vmin[ci], mx = p.planes.vmax[ci];
        if (mn < p.acc.vmin[ai]) p.acc.vmin[ai] = mn;
        if (mx > p.acc.vmax[ai]) p.acc.vmax[ai] = mx;
    }
}

// the stat record of one (key, tick) from the window's exact integer moments (generateAllStatsToQueue, :157-203)
__device__ __forceinline__ void kc_emit(apm_stat_record* rec, long long tick_ts, uint32_t key, double tpm_div, uint64_t cnt, uint64_t sum,
                                        uint64_t sq_lo, uint64_t sq_hi, uint32_t vmin, uint32_t vmax) {
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    double avg = nan, sd = nan;
    if (cnt) {
        avg = __ddiv_rn((double)sum, (double)cnt);  // :181
        const uint64_t a_lo = cnt * sq_lo, a_hi = __umul64hi(cnt, sq_lo) + cnt * sq_hi;
        const uint64_t b_lo = sum * sum, b_hi = __umul64hi(sum, sum);
        const uint64_t d_lo = a_lo - b_lo, d_hi = a_hi - b_hi - (a_lo < b_lo ? 1u : 0u);
        const double num = __dadd_rn(__dmul_rn((double)d_hi, 18446744073709551616.0), (double)d_lo);
        sd = __ddiv_rn(__dsqrt_rn(num), (double)cnt);
    }
    apm_stat_record r;
    r.tick_ts = tick_ts;
    r.key = key;
    r.flags = 1u | (cnt ? 2u : 0u);
    r.count = cnt;
    r.sum = sum;
    r.tpm = __ddiv_rn((double)cnt, tpm_div);  // :186
    r.avg = avg;
#pragma unroll
    for (int q = 0; q < APM_MAX_PCT; ++q) r.pct[q] = nan;
    r.min = cnt ? (int32_t)vmin : 0;
    r.max = (int32_t)vmax;
    r.stddev = sd;
    *rec = r;
}

__device__ __forceinline__ void kc_add_sq(uint64_t& sq_lo, uint64_t& sq_hi, uint64_t lo_part, uint64_t hi_part) {
    const uint64_t add_lo = lo_part + (hi_part << 32);
    const uint64_t add_hi = (hi_part >> 32) + (add_lo < lo_part ? 1u : 0u);
    const uint64_t t = sq_lo + add_lo;
    sq_hi += add_hi + (t < sq_lo ? 1u : 0u);
    sq_lo = t;
}

// Few bucket slots (tumbling or short windows: S <= 4): the key's accumulators stay in registers for the whole chunk and
// each tick costs ONE round trip - the segment's plane cells of every age are loaded together, unconditionally - instead
// of the dependent count -> cells -> accumulator read-modify-write chain of the general walk below.
constexpr int KC_SMALL_S = 4;
struct KcAcc {
    uint64_t sum, sq_lo, sq_hi;
    uint32_t cnt, mn, mx;
};

__global__ void __launch_bounds__(128) chunk_ticks_small_kernel(const ChunkTickParams p) {
    const uint32_t key = blockIdx.x * blockDim.x + threadIdx.x;
    if (key >= p.K) return;
    KcAcc a[KC_SMALL_S];
#pragma unroll
    for (int sl = 0; sl < KC_SMALL_S; ++sl) {
        a[sl] = KcAcc{0, 0, 0, 0, 0xFFFFFFFFu, 0};
        if (sl < p.S) {
            const size_t ai = (size_t)sl * p.K + key;
            a[sl] = KcAcc{p.acc.sum[ai], p.acc.sq_lo[ai], p.acc.sq_hi[ai], p.acc.cnt[ai], p.acc.vmin[ai], p.acc.vmax[ai]};
        }
    }
    const bool seen_before = p.seen[key] != 0;
    const uint32_t first = p.first_seg[key];
    uint32_t unseen = 0;
    for (uint32_t i = 0; i <= p.nt; ++i) {
        const bool tail = i == p.nt;   // the transactions behind the last tick are folded, nothing is emitted
        const int slot_latest = tail ? p.slot_latest_final : p.desc[i].slot_latest_before;
        // ---- fold segment i (addData of everything that arrived before the trigger, :370)
        KcAcc c[KC_SMALL_S];
#pragma unroll
        for (int age = 0; age < KC_SMALL_S; ++age) {
            c[age] = KcAcc{0, 0, 0, 0, 0xFFFFFFFFu, 0};
            if (age <= p.keep) {
                const size_t ci = ((size_t)i * (size_t)(p.keep + 1) + age) * p.K + key;
                c[age] = KcAcc{p.planes.sum[ci], p.planes.sq_lo[ci], p.planes.sq_hi[ci], p.planes.cnt[ci], p.planes.vmin[ci], p.planes.vmax[ci]};
            }
        }
#pragma unroll
        for (int age = 0; age < KC_SMALL_S; ++age) {
            if (age <= p.keep && c[age].cnt) {
                int sl = slot_latest - age;
                if (sl < 0) sl += p.S;
#pragma unroll
                for (int s2 = 0; s2 < KC_SMALL_S; ++s2)
                    if (s2 == sl) {
                        a[s2].cnt += c[age].cnt; a[s2].sum += c[age].sum; a[s2].sq_lo += c[age].sq_lo; a[s2].sq_hi += c[age].sq_hi;
                        a[s2].mn = c[age].mn < a[s2].mn ? c[age].mn : a[s2].mn;
                        a[s2].mx = c[age].mx > a[s2].mx ? c[age].mx : a[s2].mx;
                    }
            }
        }
        if (tail) break;
        const TickDesc& d = p.desc[i];
        // ---- removeOldBuckets (:103-113)
#pragma unroll
        for (int s2 = 0; s2 < KC_SMALL_S; ++s2)
            if ((d.clear >> s2) & 1ull) a[s2] = KcAcc{0, 0, 0, 0, 0xFFFFFFFFu, 0};
        apm_stat_record* rec = p.stats ? p.stats + (size_t)i * p.K + key : nullptr;
        if (!(seen_before || first <= i)) {
            if (rec) rec->flags = 0;
            ++unseen;
            continue;
        }
        if (!rec) continue;
        uint64_t cnt = 0, sum = 0, sq_lo = 0, sq_hi = 0;
        uint32_t vmin = 0xFFFFFFFFu, vmax = 0;
        for (int w = 0; w < d.nwin; ++w) {
            const int sl = d.win[w];
#pragma unroll
            for (int s2 = 0; s2 < KC_SMALL_S; ++s2)
                if (s2 == sl && a[s2].cnt) {
                    cnt += a[s2].cnt; sum += a[s2].sum;
                    kc_add_sq(sq_lo, sq_hi, a[s2].sq_lo, a[s2].sq_hi);
                    vmin = a[s2].mn < vmin ? a[s2].mn : vmin;
                    vmax = a[s2].mx > vmax ? a[s2].mx : vmax;
                }
        }
        kc_emit(rec, d.tick_ts, key, p.tpm_div, cnt, sum, sq_lo, sq_hi, vmin, vmax);
    }
#pragma unroll
    for (int sl = 0; sl < KC_SMALL_S; ++sl)
        if (sl < p.S) {
            const size_t ai = (size_t)sl * p.K + key;
            p.acc.sum[ai] = a[sl].sum; p.acc.sq_lo[ai] = a[sl].sq_lo; p.acc.sq_hi[ai] = a[sl].sq_hi;
            p.acc.cnt[ai] = a[sl].cnt; p.acc.vmin[ai] = a[sl].mn; p.acc.vmax[ai] = a[sl].mx; p.acc.run_start[ai] = 0;
        }
    if (first != KC_NO_SEG && !seen_before) p.seen[key] = 1;
    if (unseen) atomicAdd(p.unseen_records, unseen);
}

__global__ void __launch_bounds__(128) chunk_ticks_kernel(const ChunkTickParams p) {
    const uint32_t key = blockIdx.x * blockDim.x + threadIdx.x;
    if (key >= p.K) return;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const bool seen_before = p.seen[key] != 0;
    const uint32_t first = p.first_seg[key];
    uint32_t unseen = 0;
    for (uint32_t i = 0; i < p.nt; ++i) {
        const TickDesc& d = p.desc[i];
        kc_fold(p, i, d.slot_latest_before, key);              // addData of everything that arrived before the trigger (:370)
        for (int sl = 0; sl < p.S; ++sl)                        // removeOldBuckets (:103-113): evicted slots take new labels
            if ((d.clear >> sl) & 1ull) {
                const size_t ai = (size_t)sl * p.K + key;
                p.acc.sum[ai] = 0; p.acc.sq_lo[ai] = 0; p.acc.sq_hi[ai] = 0; p.acc.cnt[ai] = 0;
                p.acc.vmin[ai] = 0xFFFFFFFFu; p.acc.vmax[ai] = 0; p.acc.run_start[ai] = 0;
            }
        apm_stat_record* rec = p.stats ? p.stats + (size_t)i * p.K + key : nullptr;
        if (!(seen_before || first <= i)) {                     // not named yet: no record (:159-160 walks known keys only)
            if (rec) rec->flags = 0;
            ++unseen;
            continue;
        }
        // generateAllStatsToQueue (:157-203): window combine on exact integers
        uint64_t cnt = 0, sum = 0, sq_lo = 0, sq_hi = 0;
        uint32_t vmin = 0xFFFFFFFFu, vmax = 0;
        for (int w = 0; w < d.nwin; ++w) {
            const size_t ai = (size_t)d.win[w] * p.K + key;
            const uint32_t c = p.acc.cnt[ai];
            if (!c) continue;
            cnt += c;
            sum += p.acc.sum[ai];
            const uint64_t lo_part = p.acc.sq_lo[ai], hi_part = p.acc.sq_hi[ai];
            const uint64_t add_lo = lo_part + (hi_part << 32);
            const uint64_t add_hi = (hi_part >> 32) + (add_lo < lo_part ? 1u : 0u);
            const uint64_t t = sq_lo + add_lo;
            sq_hi += add_hi + (t < sq_lo ? 1u : 0u);
            sq_lo = t;
            const uint32_t mn = p.acc.vmin[ai], mx = p.acc.vmax[ai];
            vmin = mn < vmin ? mn : vmin;
            vmax = mx > vmax ? mx : vmax;
        }
        if (rec) {
            double avg = nan, sd = nan;
            if (cnt) {
                avg = __ddiv_rn((double)sum, (double)cnt);  // :181
                const uint64_t a_lo = cnt * sq_lo, a_hi = __umul64hi(cnt, sq_lo) + cnt * sq_hi;
                const uint64_t b_lo = sum * sum, b_hi = __umul64hi(sum, sum);
                const uint64_t d_lo = a_lo - b_lo, d_hi = a_hi - b_hi - (a_lo < b_lo ? 1u : 0u);
                const double num = __dadd_rn(__dmul_rn((double)d_hi, 18446744073709551616.0), (double)d_lo);
                sd = __ddiv_rn(__dsqrt_rn(num), (double)cnt);
            }
            apm_stat_record r;
            r.tick_ts = d.tick_ts;
            r.key = key;
            r.flags = 1u | (cnt ? 2u : 0u);
            r.count = cnt;
            r.sum = sum;
            r.tpm = __ddiv_rn((double)cnt, p.tpm_div);  // :186
            r.avg = avg;
#pragma unroll
            for (int q = 0; q < APM_MAX_PCT; ++q) r.pct[q] = nan;
            r.min = cnt ? (int32_t)vmin : 0;
            r.max = (int32_t)vmax;
            r.stddev = sd;
            *rec = r;
        }
    }
    kc_fold(p, p.nt, p.slot_latest_final, key);                 // the transactions behind the last tick
    if (first != KC_NO_SEG && !seen_before) p.seen[key] = 1;
    if (unseen) atomicAdd(p.unseen_records, unseen);
}

}  // namespace

// Is this chunk one for the whole-chunk path?  Stats only (nothing downstream needs the device to stop at a tick),
// planes of a sane size, and segments long enough that a CTA's stretch rarely straddles one.
bool apm_stats_chunk_eligible(const apm_ctx* c, uint64_t n, uint32_t nt) {
    if (!c->stats_chunk) return false;
    if ((c->cfg.flags & APM_F_PERCENTILES) && c->cfg.num_percentiles > 0) return false;
    if ((c->cfg.flags & APM_F_ZSCORE) && c->cfg.num_lags > 0) return false;
    if (c->pt_cap) return false;
    if (n < c->stats_chunk_min_n || n / ((uint64_t)nt + 1) < c->stats_chunk_min_seg) return false;
    if (nt > c->max_ticks_per_chunk) return false;
    const size_t cells = ((size_t)nt + 1) * (size_t)(c->keep + 1) * c->K;
    return cells * 36 <= ((size_t)1 << 30);
}

apm_status apm_stats_chunk(apm_ctx* c, const apm_tx_columns& d, uint64_t n, const apm_ctx::PrePass& pp, apm_tick_result* res) {
    const uint32_t nt = pp.nt;
    const bool keep_stats = (c->cfg.flags & APM_F_KEEP_STATS) != 0;
    if (keep_stats && c->ticks_buffered + nt > c->cfg.out_ticks)
        return apm_fail(c, APM_E_WOULDBLOCK, "chunk holds more ticks than the output ring has room for: fetch results or raise out_ticks");
    cudaStream_t st = c->s_main;
    // ---- planes of this chunk
    const size_t cells = ((size_t)nt + 1) * (size_t)(c->keep + 1) * c->K;
    if (cells > c->chunk_cells) {
        CUDA_TRY(c, cudaStreamSynchronize(st));
        cudaFree(c->chunk_planes);
        c->chunk_planes = nullptr;
        c->chunk_cells = 0;
        CUDA_TRY(c, cudaMalloc(&c->chunk_planes, cells * 36));
        c->chunk_cells = cells;
    }
    if (!c->chunk_first_seg) CUDA_TRY(c, cudaMalloc((void**)&c->chunk_first_seg, (size_t)c->K * sizeof(uint32_t)));
    if (!c->chunk_desc) CUDA_TRY(c, cudaMalloc(&c->chunk_desc, (size_t)(c->max_ticks_per_chunk + 1) * sizeof(TickDesc)));
    SegPlanes pl;
    {
        char* b = (char*)c->chunk_planes;
        pl.sum = (unsigned long long*)b;
        pl.sq_lo = pl.sum + cells;
        pl.sq_hi = pl.sq_lo + cells;
        pl.cnt = (uint32_t*)(pl.sq_hi + cells);
        pl.vmin = pl.cnt + cells;
        pl.vmax = pl.vmin + cells;
    }
    {
        ProfScope prof_scope_(c, APM_PG_CLEAR);
        CUDA_TRY(c, cudaMemsetAsync(pl.sum, 0, cells * 28, st));            // sums, squares, counts
        CUDA_TRY(c, cudaMemsetAsync(pl.vmin, 0xFF, cells * 4, st));
        CUDA_TRY(c, cudaMemsetAsync(pl.vmax, 0, cells * 4, st));
        CUDA_TRY(c, cudaMemsetAsync(c->chunk_first_seg, 0xFF, (size_t)c->K * sizeof(uint32_t), st));
    }
    // (the chunk's marks are on the device in position order: apm_prepass_finish put them there)
    // ---- reduce: one launch
    {
        ProfScope prof_scope_(c, APM_PG_REDUCE);
        ChunkReduceParams p;
        p.svc = d.service_id; p.jvm = d.jvm_id; p.ts = d.end_ts_ms; p.el = d.elapsed_ms; p.n = n;
        p.bucket_ms = c->cfg.bucket_ms; p.inv_bucket = c->inv_bucket; p.latest0 = c->latest; p.keep = c->keep;
        p.num_services = c->cfg.num_services; p.num_jvms = c->cfg.num_jvms; p.K = c->K;
        p.marks = pp.d_ticks; p.nt = nt; p.planes = pl; p.first_seg = c->chunk_first_seg; p.counters = c->d_counters;
        p.tiles = (uint32_t)((n + KC_TILE - 1) / KC_TILE);
        const uintptr_t align = (uintptr_t)d.service_id | (uintptr_t)d.jvm_id | (uintptr_t)d.end_ts_ms | (uintptr_t)d.elapsed_ms;
        p.vec_ok = (align & 15) == 0;
        static unsigned long long attr_done = 0;
        if (apm_first_on_device(&attr_done)) {
            CUDA_TRY(c, cudaFuncSetAttribute(reduce_chunk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(KcSmem)));
            CUDA_TRY(c, cudaFuncSetAttribute(reduce_chunk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kc_direct_bytes(KC_DIRECT_KEYS)));
        }
        unsigned g = (unsigned)c->num_sms;
        if (g > p.tiles) g = p.tiles;
        if (c->K <= KC_DIRECT_KEYS && c->stats_chunk_direct)  // every key has its own entry: no tags, no probing
            reduce_chunk_kernel<true><<<g, KC_THREADS, kc_direct_bytes(c->K), st>>>(p);
        else
            reduce_chunk_kernel<false><<<g, KC_THREADS, sizeof(KcSmem), st>>>(p);
        c->launches += 1;
        CUDA_TRY(c, cudaGetLastError());
    }
    // ---- the host side of every tick (what do_tick keeps: labels of the slots, latestBucket, counters)
    std::vector<TickDesc> desc(nt);
    const uint32_t tslot0 = keep_stats ? c->ticks_buffered : 0;
    for (uint32_t i = 0; i < nt; ++i) {
        const int64_t new_latest = pp.h_ticks[i].label, old_latest = c->latest;
        if (new_latest <= old_latest) return apm_fail(c, APM_E_INVALID, "tick label must exceed latestBucket");
        TickDesc& t = desc[i];
        t.slot_latest_before = (int)(old_latest % c->S);
        c->latest = new_latest;
        t.clear = 0;
        const int64_t first_new = std::max(old_latest + 1, new_latest - c->S + 1);
        for (int64_t lab = first_new; lab <= new_latest; ++lab) {
            if (lab < 0) continue;
            const int slot = (int)(lab % c->S);
            t.clear |= 1ull << slot;
            c->slot_label[slot] = lab;
            c->host_count[slot] = c->host_sorted[slot] = 0;
        }
        t.nwin = 0;
        for (int64_t lab = new_latest - c->keep; lab <= new_latest - c->cfg.buffer_buckets; ++lab) {
            if (lab < 0) continue;
            const int slot = (int)(lab % c->S);
            if (c->slot_label[slot] != lab) continue;
            t.win[t.nwin++] = (unsigned char)slot;
        }
        t.tick_ts = (new_latest - c->cfg.buffer_buckets - 1) * c->cfg.bucket_ms;
        c->tick_seq++;
        c->counters.ticks++;
    }
    if (keep_stats) c->ticks_buffered += nt;
    if (nt) CUDA_TRY(c, cudaMemcpyAsync(c->chunk_desc, desc.data(), nt * sizeof(TickDesc), cudaMemcpyHostToDevice, st));  // pageable: staged at once
    {
        ProfScope prof_scope_(c, APM_PG_TICKSTATS);
        ChunkTickParams tp;
        tp.acc = c->acc; tp.planes = pl; tp.first_seg = c->chunk_first_seg; tp.seen = c->seen; tp.desc = (const TickDesc*)c->chunk_desc;
        tp.nt = nt; tp.K = c->K; tp.keep = c->keep; tp.S = c->S;
        tp.slot_latest_final = (int)(c->latest % c->S);
        tp.tpm_div = c->tpm_div;
        tp.stats = keep_stats ? c->d_stats + (size_t)tslot0 * c->K : nullptr;
        tp.unseen_records = &c->d_counters->unseen_records;
        if (c->S <= KC_SMALL_S) chunk_ticks_small_kernel<<<(c->K + 127) / 128, 128, 0, st>>>(tp);
        else chunk_ticks_kernel<<<(c->K + 127) / 128, 128, 0, st>>>(tp);
        c->launches += 1;
        CUDA_TRY(c, cudaGetLastError());
    }
    if (res) res->ticks += nt;
    return APM_OK;
}

cudaError_t apm_preload_k1_chunk() {
    const void* fns[] = {(const void*)reduce_chunk_kernel<false>, (const void*)reduce_chunk_kernel<true>, (const void*)chunk_ticks_kernel, (const void*)chunk_ticks_small_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
