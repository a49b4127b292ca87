// K4: rolling baseline, signal and lag-list push.
//
// Replaces processData + processZScoreStats (stream_calc_z_score.js:195-311, 66-104) and
// the Array.prototype average / standardDeviation they call (util_methods.js:10-50) for all
// keys of one lag at once.
//
// Layout: ring[slot][series][key] (time-major), one thread per (series, key) column, so a
// warp reads 32 consecutive doubles per slot.  Every list of a lag advances in lock-step
// (each seen key gets exactly one StatEntry per tick), therefore one global head per lag
// serves all keys: a list of length n occupies the n slots before `head`, the push goes to
// `head`, and a full list is summed head, head+1, ... (oldest first) - the same left-to-right
// order as Array.prototype.reduce, so the fp64 mean has the same bits as the reference's.
#include <cstring>
#include <string>

#include "ctx.h"

namespace {

constexpr int K4_THREADS = 128;
constexpr int K4_UNROLL = 32;

struct BaselineParams {
    double* ring;          // [L][3K]
    uint32_t* ring_len;    // [3K]
    uint32_t* ring_next;   // [3K] slot each column's next push belongs in (== head for every lock-step list)
    const double* zin;     // [4][K]: tpm', val0..2
    const double* thr;     // [K]
    const double* infl;    // [K]
    const uint8_t* seen;   // [K] snapshot taken by the tick's window kernel (emit flags)
    apm_fs_record* fs;     // [K] records of this (tick, lag); nullptr when not retained
    uint32_t K, L, head, lag_value;
    uint32_t pitch;        // doubles per ring row
    int64_t tick_ts;
    int sd_mode;
};

__device__ __forceinline__ void accumulate(double v, double& sum, uint32_t& cnt) {
    // average(): `value != null && !isNaN(parseFloat(value))` then `sum + parseFloat(value)`.
    // A skipped entry is added as +0.0 instead of branching around the add: the running sum starts
    // at +0.0 and can therefore never be -0.0 (x + y is -0.0 only when both are), so s + 0.0 == s
    // bit for bit, and the select sits on the loaded value - off the dependent chain of adds, which
    // is what bounds this kernel when only a few hundred lists are walked.
    const bool ok = v == v;
    sum = __dadd_rn(sum, ok ? v : 0.0);
    cnt += ok ? 1u : 0u;
}

__global__ void __launch_bounds__(K4_THREADS) baseline_kernel(const BaselineParams p) {
    const uint32_t col = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t ncol = p.pitch;
    if (col >= 3u * p.K) return;
    const uint32_t series = col / p.K, key = col - series * p.K;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    apm_fs_record* rec = p.fs ? p.fs + key : nullptr;
    if (!p.seen[key]) {
        if (rec && series == 0) rec->flags = 0;
        return;
    }
    const uint32_t len = p.ring_len[col];
    const double new_value = p.zin[(size_t)(1 + series) * p.K + key];
    const double threshold = p.thr[key], influence = p.infl[key];
    double infl_value = new_value, avg = nan, lb = nan, ub = nan;
    int signal = 0;

    if (len >= p.L) {  // prevValuesList.length >= lag (:75)
        double sum = 0.0, sumsq_dev = 0.0;
        uint32_t cnt = 0;
        const double* base = p.ring + col;
        // Oldest first: element i of the list lives in slot (head + i) mod L.  The chain of
        // additions is sequential by definition, so bandwidth comes from loads in flight: two
        // register batches are software-pipelined (one being added while the other is loading).
        // Reads past the end are replaced by NaN, which average() skips anyway.
        auto load_elem = [&](uint32_t i) -> double {
            uint32_t sl = p.head + i;
            sl = sl >= p.L ? sl - p.L : sl;
            return i < p.L ? __ldcs(base + (size_t)sl * ncol) : nan;
        };
        double va[K4_UNROLL], vb[K4_UNROLL];
#pragma unroll
        for (int u = 0; u < K4_UNROLL; ++u) va[u] = load_elem(u);
        for (uint32_t i = 0; i < p.L; i += 2 * K4_UNROLL) {
#pragma unroll
            for (int u = 0; u < K4_UNROLL; ++u) vb[u] = load_elem(i + K4_UNROLL + u);
#pragma unroll
            for (int u = 0; u < K4_UNROLL; ++u) accumulate(va[u], sum, cnt);
#pragma unroll
            for (int u = 0; u < K4_UNROLL; ++u) va[u] = load_elem(i + 2 * K4_UNROLL + u);
#pragma unroll
            for (int u = 0; u < K4_UNROLL; ++u) accumulate(vb[u], sum, cnt);
        }
        const double last = base[(size_t)((p.head + p.L - 1) % p.L) * ncol];
        bool avg_ok = cnt > 0, sd_ok = false;
        double sd = nan;
        if (avg_ok) {
            avg = __ddiv_rn(sum, (double)cnt);
            if (p.sd_mode == APM_SD_REFERENCE) {
                // standardDeviation(): average(squareDiffs) ignores its argument -> sqrt(mean);
                // mean == 0 -> undefined, mean < 0 -> NaN (util_methods.js:41-49)
                if (avg != 0.0) sd = __dsqrt_rn(avg);
                sd_ok = sd == sd;
            } else {
                // extension: population stddev of the valid entries, second exact-order pass
                for (int span = 0; span < 2; ++span) {
                    uint32_t s = span == 0 ? p.head : 0u;
                    const uint32_t e = span == 0 ? p.L : p.head;
                    for (; s < e; ++s) {
                        const double v = base[(size_t)s * ncol];
                        if (v == v) {
                            const double d = __dsub_rn(v, avg);
                            sumsq_dev = __dadd_rn(sumsq_dev, __dmul_rn(d, d));
                        }
                    }
                }
                sd = __dsqrt_rn(__ddiv_rn(sumsq_dev, (double)cnt));
                sd_ok = sd == sd;
            }
        }
        if (avg_ok && sd_ok) {  // (:80-83) no fused multiply-add: V8 rounds the product first
            const double band = __dmul_rn(threshold, sd);
            lb = __dsub_rn(avg, band);
            ub = __dadd_rn(avg, band);
            if (new_value == new_value) {  // (:87) NaN new value never signals
                if (fabs(__dsub_rn(new_value, avg)) > band) {  // strict (:89)
                    signal = new_value > avg ? 1 : -1;
                    if (last == last)  // (:96-97)
                        infl_value = __dadd_rn(__dmul_rn(influence, new_value),
                                               __dmul_rn(__dsub_rn(1.0, influence), last));
                }
            }
        }
    } else {
        p.ring_len[col] = len + 1;
    }
    // shift() + push() (:285-306): the oldest slot is the one the push overwrites
    p.ring[(size_t)p.head * ncol + col] = infl_value;
    p.ring_next[col] = p.head + 1 == p.L ? 0u : p.head + 1;

    if (rec) {
        if (series == 0) {
            rec->tick_ts = p.tick_ts;
            rec->key = key;
            rec->lag = p.lag_value;
            rec->tpm = p.zin[key];
            rec->flags = 1;
            rec->pad_[0] = rec->pad_[1] = rec->pad_[2] = rec->pad_[3] = 0;
        }
        rec->val[series] = new_value;
        rec->avg[series] = avg;
        rec->lb[series] = lb;
        rec->ub[series] = ub;
        rec->signal[series] = (int8_t)signal;
    }
}

// ---- the same stage with the ring streamed by TMA --------------------------------------------
// One warp per CTA owns 32 adjacent columns.  Lane 0 keeps K4T_STAGES bulk-tensor loads
// (cp.async.bulk.tensor.2d, box = 32 columns x K4T_ROWS slots = 4 KB) in flight into shared
// memory, each completing on its own mbarrier; the lanes then add their column's values in
// slot order out of shared memory.  The additions of a column are sequential by definition
// (that is what makes the mean bit-identical to the reference's), so all the bandwidth comes
// from the depth of the load pipeline, which TMA provides without spending registers.
constexpr int K4T_COLS = 32;
constexpr int K4T_ROWS = 16;
constexpr uint32_t K4T_STAGE_BYTES = K4T_COLS * K4T_ROWS * sizeof(double);
constexpr int K4T_DEEP_STAGES = 8;   // x 3 boxes of 4 KB = the same 96 KB in flight per CTA

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// The ring is read exactly once per tick and is far larger than L2: mark its lines evict-first so
// the stream does not push the ingest / sort working set (a few tens of MB, re-read by the next
// kernel of that chain) out of L2 while both streams run.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int r0, uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;"
        ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(r0), "r"(smem_u32(bar)), "l"(policy) : "memory");
}

// What follows the walk of a column: mean, "stddev", bounds, signal, influence blend, push, fs record
// (processZScoreStats :75-104 and the shift() + push() of processData :285-306).
__device__ __forceinline__ void finish_column(const BaselineParams& p, uint32_t col, uint32_t series, uint32_t key, bool seen, bool full,
                                              uint32_t len, double sum, uint32_t cnt, double last) {
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    apm_fs_record* rec = p.fs ? p.fs + key : nullptr;
    if (!seen) {
        if (rec && series == 0) rec->flags = 0;
        return;
    }
    const double new_value = p.zin[(size_t)(1 + series) * p.K + key];
    const double threshold = p.thr[key], influence = p.infl[key];
    double infl_value = new_value, avg = nan, lb = nan, ub = nan;
    int signal = 0;
    if (full) {
        if (cnt > 0) {
            avg = __ddiv_rn(sum, (double)cnt);
            // standardDeviation() == sqrt(mean): mean == 0 -> undefined, mean < 0 -> NaN (util_methods.js:41-49)
            double sd = nan;
            if (avg != 0.0) sd = __dsqrt_rn(avg);
            if (sd == sd) {
                const double band = __dmul_rn(threshold, sd);  // (:80-83), product rounded before the add
                lb = __dsub_rn(avg, band);
                ub = __dadd_rn(avg, band);
                if (new_value == new_value && fabs(__dsub_rn(new_value, avg)) > band) {  // (:87-89) strict
                    signal = new_value > avg ? 1 : -1;
                    if (last == last)  // (:96-97)
                        infl_value = __dadd_rn(__dmul_rn(influence, new_value), __dmul_rn(__dsub_rn(1.0, influence), last));
                }
            }
        }
    } else {
        p.ring_len[col] = len + 1;
    }
    p.ring[(size_t)p.head * p.pitch + col] = infl_value;  // shift() + push() (:285-306)
    p.ring_next[col] = p.head + 1 == p.L ? 0u : p.head + 1;
    if (rec) {
        if (series == 0) {
            rec->tick_ts = p.tick_ts;
            rec->key = key;
            rec->lag = p.lag_value;
            rec->tpm = p.zin[key];
            rec->flags = 1;
            rec->pad_[0] = rec->pad_[1] = rec->pad_[2] = rec->pad_[3] = 0;
        }
        rec->val[series] = new_value;
        rec->avg[series] = avg;
        rec->lb[series] = lb;
        rec->ub[series] = ub;
        rec->signal[series] = (int8_t)signal;
    }
}

// MULT: 16-slot boxes per stage.  The bandwidth-bound instance (many CTAs per SM) uses one; the instance for a
// few hundred lists, where one warp per SM is paced by per-chunk overhead, uses three.
template <int K4T_STAGES, int MULT>
struct __align__(128) K4TSmem {
    double buf[K4T_STAGES][K4T_ROWS * MULT][K4T_COLS];
    uint64_t bar[K4T_STAGES];
};

template <int K4T_STAGES, int MULT>
__global__ void __launch_bounds__(32) baseline_tma_kernel(const __grid_constant__ CUtensorMap tmap, const BaselineParams p) {
    constexpr int ROWS = K4T_ROWS * MULT;
    // dynamic shared memory: the deep-pipeline instance (few columns, long lists) needs more than 48 KB
    // (no static shared memory in this kernel, so the dynamic window starts at offset 0 and is 128-byte aligned)
    extern __shared__ __align__(128) unsigned char k4t_smem_raw[];
    K4TSmem<K4T_STAGES, MULT>& sm = *reinterpret_cast<K4TSmem<K4T_STAGES, MULT>*>(k4t_smem_raw);
    const int lane = threadIdx.x;
    const uint32_t col0 = blockIdx.x * K4T_COLS, col = col0 + lane;
    const bool in_range = col < 3u * p.K;
    const uint32_t series = in_range ? col / p.K : 0, key = in_range ? col - series * p.K : 0;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const bool seen = in_range && p.seen[key];
    const uint32_t len = seen ? p.ring_len[col] : 0;
    const bool full = seen && len >= p.L;  // prevValuesList.length >= lag (:75)

    double sum = 0.0, last = nan;
    uint32_t cnt = 0;
    if (__any_sync(0xffffffffu, full)) {
        const uint64_t policy = l2_evict_first_policy();
        // oldest first: rows head .. L-1, then 0 .. head-1, in chunks of ROWS slots
        const uint32_t nA = (p.L - p.head + ROWS - 1) / ROWS, nB = (p.head + ROWS - 1) / ROWS;
        const uint32_t n = nA + nB;
        auto row0_of = [&](uint32_t c) { return c < nA ? p.head + c * ROWS : (c - nA) * ROWS; };
        auto rows_of = [&](uint32_t c) {
            const uint32_t r0 = row0_of(c), end = c < nA ? p.L : p.head;
            return end - r0 < (uint32_t)ROWS ? end - r0 : (uint32_t)ROWS;
        };
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < K4T_STAGES; ++s) mbar_init(&sm.bar[s], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        __syncwarp();
        if (lane == 0) {
            for (uint32_t c = 0; c < n && c < (uint32_t)K4T_STAGES; ++c) {
                mbar_expect_tx(&sm.bar[c], K4T_STAGE_BYTES * MULT);  // rows past the end are zero-filled but still counted
#pragma unroll
                for (int m = 0; m < MULT; ++m)
                    tma_load_2d(&sm.buf[c][m * (ROWS / MULT)][0], &tmap, (int)col0, (int)row0_of(c) + m * (ROWS / MULT), &sm.bar[c], policy);
            }
        }
        // Software pipeline over the chunks: the values of chunk c + 1 are pulled from shared memory
        // into registers before the dependent adds of chunk c run, so with one warp per SM (few
        // lists) the walk is paced by the add chain, not by barrier and shared-memory latency.
        // The two register sets swap roles from one chunk to the next (the loop advances two chunks
        // per trip): copying one set into the other cost 2.3 of the 6.1 instructions per slot, and a
        // lone warp issues one instruction every four to five cycles.
        auto step = [&](uint32_t c, double (&v)[ROWS], double (&vn)[ROWS]) {
            const uint32_t s = c % K4T_STAGES;
            if (c + 1 < n) {
                const uint32_t s1 = (c + 1) % K4T_STAGES, parity1 = ((c + 1) / K4T_STAGES) & 1u;
                mbar_wait(&sm.bar[s1], parity1);
#pragma unroll
                for (int r = 0; r < ROWS; ++r) vn[r] = sm.buf[s1][r][lane];
            }
            const uint32_t rows = rows_of(c);
            if (rows == (uint32_t)ROWS) {
                if (full) {
                    // Bare adds, validity checked once on the result: a NaN entry (a tick
                    // the key had no data for) poisons t, and only then is the chunk redone entry
                    // by entry from the saved sum.  Without a NaN the adds are the same sequence.
                    double t = sum;
#pragma unroll
                    for (int r = 0; r < ROWS; ++r) t = __dadd_rn(t, v[r]);
                    if (t == t) {
                        sum = t;
                        cnt += ROWS;
                    } else {
#pragma unroll
                        for (int r = 0; r < ROWS; ++r) accumulate(v[r], sum, cnt);
                    }
                }
                last = v[ROWS - 1];
            } else {
#pragma unroll
                for (int r = 0; r < ROWS; ++r)
                    if ((uint32_t)r < rows) {
                        if (full) accumulate(v[r], sum, cnt);
                        last = v[r];
                    }
            }
            __syncwarp();  // every lane has consumed chunk c from registers: its stage may be overwritten
            if (lane == 0 && c + K4T_STAGES < n) {
                mbar_expect_tx(&sm.bar[s], K4T_STAGE_BYTES * MULT);
#pragma unroll
                for (int m = 0; m < MULT; ++m)
                    tma_load_2d(&sm.buf[s][m * (ROWS / MULT)][0], &tmap, (int)col0, (int)row0_of(c + K4T_STAGES) + m * (ROWS / MULT), &sm.bar[s], policy);
            }
        };
        double va[ROWS], vb[ROWS];
        mbar_wait(&sm.bar[0], 0);
#pragma unroll
        for (int r = 0; r < ROWS; ++r) va[r] = sm.buf[0][r][lane];
        for (uint32_t c = 0; c < n; c += 2) {
            step(c, va, vb);
            if (c + 1 < n) step(c + 1, vb, va);
        }
    }
    if (!in_range) return;
    finish_column(p, col, series, key, seen, full, len, sum, cnt, last);
}

// ---- the walk for a few hundred lists (the reference's own deployment; small key shards) -------------------------
// With fewer column groups than SMs every warp has an SM to itself and the walk is no longer a bandwidth problem: it
// is one chain of dependent DADDs per lane, 8 cycles each (scripts/micro_dadd.cu), i.e. L x 8 cycles whatever else
// happens.  A single warp that also waits on barriers, moves shared memory to registers and hands stages back spends
// 20-odd cycles per slot (ncu: the 48 LDS of a chunk issue in front of its 48 DADDs, not under them, and everything
// between two chunks is serial too).  So the jobs get a warp each:
//   warp 0, one elected lane: waits for a stage to be empty, arms its `full` barrier, issues ONE bulk-tensor load of
//           96 slots x 32 columns (24 KB);
//   warps 1 and 2 take the chunks in turn.  While one runs its chunk's 96 dependent adds - nothing else in its
//           instruction stream - the other has already pulled the next chunk into registers, checked it for
//           undefined entries and waits at a named barrier for the running {sum, count}, which crosses in a 16-byte
//           mailbox in shared memory; the hand-over costs one barrier wake-up and one shared-memory load per 96 slots
//           (scripts/micro_k4few.cu: 897 cycles per chunk against 772 for the bare chain; 965 with a polled mailbox).
// All lags go into ONE launch (CTA ranges per lag, the longest list first): the short walks run beside the long one on
// other SMs instead of in front of it.
struct alignas(64) FewParams {
    CUtensorMap tmap[APM_MAX_LAGS];
    BaselineParams p[APM_MAX_LAGS];
    uint32_t cta_end[APM_MAX_LAGS];  // launch order: CTAs [cta_end[i - 1], cta_end[i]) walk lag lag_of[i]
    int32_t lag_of[APM_MAX_LAGS];
    int32_t n;
    // alert compaction by the last CTA to finish (compact_n = 0: a separate launch does it)
    uint32_t compact_n;              // records of this tick, [lag][key]
    uint32_t alert_cap;
    const apm_fs_record* fs_tick;
    apm_fs_record* alerts;
    unsigned int* alert_count;
    uint32_t* done;                  // CTAs that have written their records; left at 0
};

constexpr int K4F_ROWS = 96;   // slots per chunk = per bulk load (box 32 x 96 doubles)
constexpr int K4F_STAGES = 6;  // two chunks sit in the adders' registers with their stages still held: four in flight
constexpr uint32_t K4F_STAGE_BYTES = K4T_COLS * K4F_ROWS * sizeof(double);

struct alignas(16) K4FMail {
    double sum;
    uint32_t cnt;
    uint32_t pad;
};

struct __align__(128) K4FSmem {
    double buf[K4F_STAGES][K4F_ROWS][K4T_COLS];
    uint64_t full[K4F_STAGES];
    uint64_t empty[K4F_STAGES];
    K4FMail mail[2][32];  // [written by adder w][lane]
};

// The signalled records of one tick, in record order, into the tick's alert block: compact_alerts_small_kernel's
// result by ONE warp (each lane owns a contiguous stretch of records, so the order survives a lane-level prefix sum).
// fs was written by other CTAs: read through L2.
__device__ void compact_by_warp(const apm_fs_record* fs, uint32_t n, apm_fs_record* alerts, unsigned int* count, uint32_t cap, int lane) {
    // n <= 4096: a lane owns at most 128 records, whose hits it keeps as a bit mask between the two passes
    const uint32_t per = (n + 31u) / 32u, lo = (uint32_t)lane * per, hi = lo + per < n ? lo + per : n;
    uint32_t mask[4] = {0u, 0u, 0u, 0u};
    for (uint32_t b0 = 0; b0 < per; b0 += 16) {
        uint32_t w[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) {  // bytes 120..122: the three signals, byte 123: flags (bit 0 = record present)
            const uint32_t i = lo + b0 + u;
            w[u] = i < hi ? __ldcg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(fs + i) + 120)) : 0u;
        }
        uint32_t bits = 0;
#pragma unroll
        for (int u = 0; u < 16; ++u) bits |= ((w[u] & 0x01000000u) && (w[u] & 0x00ffffffu)) ? 1u << u : 0u;
        mask[b0 >> 5] |= bits << (b0 & 31u);
    }
    const uint32_t mine = __popc(mask[0]) + __popc(mask[1]) + __popc(mask[2]) + __popc(mask[3]);
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t other = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += other;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t k = incl - mine;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint32_t bits = mask[q];
        while (bits) {
            const uint32_t i = lo + 32u * q + (uint32_t)__ffs((int)bits) - 1u;
            bits &= bits - 1u;
            if (k < cap) {
                const uint4* src = reinterpret_cast<const uint4*>(fs + i);
                uint4* dst = reinterpret_cast<uint4*>(alerts + k);
                uint4 r[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) r[j] = __ldcg(src + j);
#pragma unroll
                for (int j = 0; j < 8; ++j) dst[j] = r[j];
            }
            ++k;
        }
    }
    if (lane == 0) *count = total;
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

#ifdef APM_K4_DEBUG
__device__ unsigned long long apm_k4_dbg[2][16];
#define K4DBG(i) do { if (dbg) { const long long t_ = clock64(); acc[i] += t_ - tprev; tprev = t_; } } while (0)
#else
#define K4DBG(i) do { } while (0)
#endif

__global__ void __launch_bounds__(96) baseline_few_kernel(const __grid_constant__ FewParams fp) {
#ifdef APM_K4_DEBUG
    const long long t_entry = clock64();
    long long acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = t_entry;
#endif
    constexpr int ROWS = K4F_ROWS, STAGES = K4F_STAGES;
    extern __shared__ __align__(128) unsigned char k4f_smem_raw[];
    K4FSmem& sm = *reinterpret_cast<K4FSmem*>(k4f_smem_raw);
    int slot = 0;
    while (slot + 1 < fp.n && blockIdx.x >= fp.cta_end[slot]) ++slot;
    const int l = fp.lag_of[slot];
    const BaselineParams& p = fp.p[l];
    const CUtensorMap* tmap = &fp.tmap[l];
    const uint32_t cta = blockIdx.x - (slot ? fp.cta_end[slot - 1] : 0u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t L = p.L, head = p.head;
    const uint32_t col0 = cta * K4T_COLS, col = col0 + lane;
    const bool in_range = col < 3u * p.K;
    const uint32_t series = in_range ? col / p.K : 0, key = in_range ? col - series * p.K : 0;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const bool seen = in_range && p.seen[key];
    const uint32_t len = seen ? p.ring_len[col] : 0;
    const bool full = seen && len >= L;  // prevValuesList.length >= lag (:75)
    const bool any_full = __any_sync(0xffffffffu, full);  // every warp looks at the same 32 columns: the same answer

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    // oldest first: rows head .. L-1, then 0 .. head-1, in chunks of ROWS slots
    const uint32_t nA = (L - head + ROWS - 1) / ROWS, nB = (head + ROWS - 1) / ROWS;
    const uint32_t n = any_full ? nA + nB : 0u;

    if (warp == 0) {  // ---- the loader
        if (lane != 0) return;
        const uint64_t policy = l2_evict_first_policy();
        for (uint32_t c = 0; c < n; ++c) {
            const uint32_t s = c % STAGES;
            if (c >= (uint32_t)STAGES) mbar_wait(&sm.empty[s], ((c / STAGES) - 1u) & 1u);
            mbar_expect_tx(&sm.full[s], K4F_STAGE_BYTES);  // rows past the end are zero-filled but still counted
            const uint32_t r0 = c < nA ? head + c * ROWS : (c - nA) * ROWS;
            tma_load_2d(&sm.buf[s][0][0], tmap, (int)col0, (int)r0, &sm.full[s], policy);
        }
        return;
    }

    // ---- the adders: w takes chunks w, w + 2, ...
    const uint32_t w = (uint32_t)warp - 1u;
#ifdef APM_K4_DEBUG
    const bool dbg = blockIdx.x == 0 && lane == 0;
    K4DBG(0);  // prologue
#endif
    double sum = 0.0, last = nan;
    uint32_t cnt = 0;
    for (uint32_t c = w; c < n; c += 2) {
        const uint32_t s = c % STAGES;
        mbar_wait(&sm.full[s], (c / STAGES) & 1u);
        K4DBG(1);  // waiting for the stage
        double v[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) v[r] = sm.buf[s][r][lane];
#ifdef APM_K4_DEBUG
        if (dbg && __double_as_longlong(v[ROWS - 1]) == 0x123456789abcdefll) acc[7]++;  // the loads return in order: wait for the last
#endif
        K4DBG(2);  // shared memory -> registers
        const uint32_t r0 = c < nA ? head + c * ROWS : (c - nA) * ROWS, end = c < nA ? L : head;
        const uint32_t rows = end - r0 < (uint32_t)ROWS ? end - r0 : (uint32_t)ROWS;
        // Still under the other adder's chain: the chunk is made safe for BARE adds.  average() skips an undefined entry
        // (util_methods.js:12-18); adding +0.0 in its place gives the same bits, because the running sum starts at +0.0
        // and can never be -0.0 - so a NaN entry, and a row past the end of a ragged chunk, becomes +0.0 here and is
        // left out of the count.  (Whether anything needs replacing is first read off the exponent fields: one integer
        // max per entry.)  The chain below is then the same 96 adds for every chunk, and its result can be posted the
        // moment the last one retires.
        uint32_t hi_max = 0;
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            const uint32_t hi = (uint32_t)__double2hiint(v[r]) & 0x7fffffffu;
            hi_max = hi > hi_max ? hi : hi_max;
        }
        uint32_t nvalid = ROWS;
        last = sm.buf[s][rows - 1u][lane];  // prevValuesList[length - 1], undefined or not (:96); read from the stage: a
                                            // run-time index into v[] would move the whole array to local memory
        if (__any_sync(0xffffffffu, hi_max >= 0x7ff00000u) || rows != (uint32_t)ROWS) {
            nvalid = 0;
#pragma unroll
            for (int r = 0; r < ROWS; ++r) {
                const bool here = (uint32_t)r < rows;
                const bool ok = here && v[r] == v[r];
                nvalid += ok ? 1u : 0u;
                v[r] = ok ? v[r] : 0.0;
            }
        }
        if (c > 0) {  // the running sum after chunk c - 1, from the other adder
            asm volatile("bar.sync %0, 64;" ::"r"(1u + (w ^ 1u)) : "memory");
            const uint32_t addr = smem_u32(&sm.mail[w ^ 1u][lane]);
            unsigned long long a, b;
            asm volatile("ld.volatile.shared.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr) : "memory");
            sum = __longlong_as_double((long long)a);
            cnt = (uint32_t)b;
        }
        K4DBG(3);  // waiting for the other adder
        // (a lane whose list is not full yet computes a sum nobody reads)
#pragma unroll
        for (int r = 0; r < ROWS; ++r) sum = __dadd_rn(sum, v[r]);
        cnt += nvalid;
        if (c + 1 < n) {
            const uint32_t addr = smem_u32(&sm.mail[w][lane]);
            const unsigned long long a = (unsigned long long)__double_as_longlong(sum), b = (unsigned long long)cnt;
            asm volatile("st.volatile.shared.v2.b64 [%0], {%1, %2};" ::"r"(addr), "l"(a), "l"(b) : "memory");
            asm volatile("bar.arrive %0, 64;" ::"r"(1u + w) : "memory");
        }
        K4DBG(4);  // the chain
        // the chunk's values have all been used: its stage may be overwritten
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[s]);
        K4DBG(5);  // post + release
    }
#ifdef APM_K4_DEBUG
    if (dbg) {
        for (int i = 0; i < 8; ++i) apm_k4_dbg[w][i] = (unsigned long long)acc[i];
        apm_k4_dbg[w][8] = (unsigned long long)(clock64() - t_entry);
        apm_k4_dbg[w][9] = n;
    }
#endif
    // the adder that folded the last chunk holds the result (adder 0 when nothing was walked)
    if (((n ? n - 1u : 0u) & 1u) != w) return;
    if (in_range) finish_column(p, col, series, key, seen, full, len, sum, cnt, last);
    if (fp.compact_n) {
        // one finishing warp per CTA: the one that draws the last ticket sees every CTA's records
        __syncwarp();
        __threadfence();
        uint32_t ticket = 0;
        if (lane == 0) ticket = atomicAdd(fp.done, 1u);
        ticket = __shfl_sync(0xffffffffu, ticket, 0);
        if (ticket == gridDim.x - 1u) {
            if (lane == 0) *fp.done = 0;  // for the next launch (stream-ordered behind this one)
            __threadfence();
            compact_by_warp(fp.fs_tick, fp.compact_n, fp.alerts, fp.alert_count, fp.alert_cap, lane);
        }
    }
}

// ---- guarded incremental baseline (APM_F_INCREMENTAL) ------------------------------------------
// The exact-order rescan above costs K x L x 24 B per tick.  This variant keeps, per column, a
// compensated running sum of the list (Neumaier), the number of valid entries and the number of
// non-zero entries, updated with the evicted and the pushed value: O(K) per tick.  Its mean
// differs from the reference's left-to-right sum by rounding only (both are within ~L x 2^-53
// relative of the true mean), so the record's mean / bounds agree to ~1e-12 relative (the spec
// allows 1e-9) - but a signal must be bit-identical.  Therefore the strict test
// |new - mean| > threshold * sqrt(mean) is decided from the running mean only when its margin is
// 1e-9 relative or more; inside that guard band the thread re-sums its column in the reference's
// order and decides exactly as the exact kernel does.  Pushed values depend on the signal and on
// ring contents only, so the lists stay bit-identical to the reference's.
// Assumes non-negative series (durations); a non-positive running mean over non-zero entries also
// takes the exact path.
struct IncState {
    double* sum;      // [3K]
    double* comp;     // [3K] Neumaier compensation
    uint32_t* cnt;    // [3K] valid (non-NaN) entries
    uint32_t* nz;     // [3K] valid entries != 0
};

__device__ __forceinline__ void neumaier_add(double& s, double& c, double x) {
    const double t = __dadd_rn(s, x);
    if (fabs(s) >= fabs(x)) c = __dadd_rn(c, __dadd_rn(__dsub_rn(s, t), x));
    else c = __dadd_rn(c, __dadd_rn(__dsub_rn(x, t), s));
    s = t;
}

// left-to-right sum of a full list in the reference's order (slots head, head+1, ...)
__device__ void exact_column_sum(const double* base, uint32_t head, uint32_t L, size_t pitch, double& sum, uint32_t& cnt) {
    sum = 0.0;
    cnt = 0;
    uint32_t i = 0;
    for (; i + 8 <= L; i += 8) {
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            uint32_t sl = head + i + u;
            sl = sl >= L ? sl - L : sl;
            v[u] = base[(size_t)sl * pitch];
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) accumulate(v[u], sum, cnt);
    }
    for (; i < L; ++i) {
        uint32_t sl = head + i;
        sl = sl >= L ? sl - L : sl;
        accumulate(base[(size_t)sl * pitch], sum, cnt);
    }
}

__global__ void __launch_bounds__(K4_THREADS) baseline_inc_kernel(const BaselineParams p, const IncState st,
                                                                  unsigned long long* __restrict__ fallbacks) {
    const uint32_t col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= 3u * p.K) return;
    const uint32_t series = col / p.K, key = col - series * p.K;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    apm_fs_record* rec = p.fs ? p.fs + key : nullptr;
    if (!p.seen[key]) {
        if (rec && series == 0) rec->flags = 0;
        return;
    }
    const uint32_t len = p.ring_len[col];
    const double new_value = p.zin[(size_t)(1 + series) * p.K + key];
    const double threshold = p.thr[key], influence = p.infl[key];
    double infl_value = new_value, avg = nan, lb = nan, ub = nan;
    int signal = 0;
    double S = st.sum[col], C = st.comp[col];
    uint32_t cnt = st.cnt[col], nz = st.nz[col];
    double* base = p.ring + col;
    const bool full = len >= p.L;
    double evicted = nan;
    if (full) {
        evicted = base[(size_t)p.head * p.pitch];
        const double last = base[(size_t)(p.head == 0 ? p.L - 1 : p.head - 1) * p.pitch];
        if (cnt > 0) {
            bool exact = false;
            double m = __ddiv_rn(__dadd_rn(S, C), (double)cnt);
            if (nz == 0) m = 0.0;                    // all valid entries are zero: the reference's mean is exactly 0
            else if (!(m > 0.0)) exact = true;       // negative entries in play: no shortcut
            double sd = nan, band = 0.0;
            if (!exact && m != 0.0) {
                sd = __dsqrt_rn(m);
                band = __dmul_rn(threshold, sd);
                if (new_value == new_value) {
                    const double d = fabs(__dsub_rn(new_value, m)) - band;
                    if (fabs(d) <= 1e-9 * (fabs(new_value) + m + fabs(band))) exact = true;  // inside the guard band
                }
            }
            if (exact) {
                double es;
                uint32_t ec;
                exact_column_sum(base, p.head, p.L, p.pitch, es, ec);
                m = __ddiv_rn(es, (double)ec);
                sd = nan;
                if (m != 0.0) sd = __dsqrt_rn(m);
                band = __dmul_rn(threshold, sd);
                atomicAdd(fallbacks, 1ull);
            }
            avg = m;
            if (sd == sd) {
                lb = __dsub_rn(avg, band);
                ub = __dadd_rn(avg, band);
                if (new_value == new_value && fabs(__dsub_rn(new_value, avg)) > band) {
                    signal = new_value > avg ? 1 : -1;
                    if (last == last)
                        infl_value = __dadd_rn(__dmul_rn(influence, new_value), __dmul_rn(__dsub_rn(1.0, influence), last));
                }
            }
        }
    } else {
        p.ring_len[col] = len + 1;
    }
    // running statistics follow the shift() + push()
    if (evicted == evicted) {
        neumaier_add(S, C, -evicted);
        cnt -= 1;
        if (evicted != 0.0) nz -= 1;
    }
    if (infl_value == infl_value) {
        neumaier_add(S, C, infl_value);
        cnt += 1;
        if (infl_value != 0.0) nz += 1;
    }
    if (cnt == 0) { S = 0.0; C = 0.0; }
    st.sum[col] = S; st.comp[col] = C; st.cnt[col] = cnt; st.nz[col] = nz;
    base[(size_t)p.head * p.pitch] = infl_value;
    p.ring_next[col] = p.head + 1 == p.L ? 0u : p.head + 1;
    if (rec) {
        if (series == 0) {
            rec->tick_ts = p.tick_ts;
            rec->key = key;
            rec->lag = p.lag_value;
            rec->tpm = p.zin[key];
            rec->flags = 1;
            rec->pad_[0] = rec->pad_[1] = rec->pad_[2] = rec->pad_[3] = 0;
        }
        rec->val[series] = new_value;
        rec->avg[series] = avg;
        rec->lb[series] = lb;
        rec->ub[series] = ub;
        rec->signal[series] = (int8_t)signal;
    }
}

// (Re)builds the running statistics of every column from its list (after lists were loaded).
__global__ void __launch_bounds__(K4_THREADS) inc_init_kernel(const double* __restrict__ ring, const uint32_t* __restrict__ ring_len,
                                                              const uint32_t* __restrict__ ring_next, const IncState st, uint32_t K,
                                                              uint32_t L, size_t pitch) {
    const uint32_t col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= 3u * K) return;
    const uint32_t len = ring_len[col] < L ? ring_len[col] : L;
    const uint32_t first = (ring_next[col] + L - len) % L;
    double S = 0.0, C = 0.0;
    uint32_t cnt = 0, nz = 0;
    for (uint32_t i = 0; i < len; ++i) {
        uint32_t sl = first + i;
        sl = sl >= L ? sl - L : sl;
        const double v = __ldcs(ring + (size_t)sl * pitch + col);
        if (v == v) {
            neumaier_add(S, C, v);
            cnt += 1;
            if (v != 0.0) nz += 1;
        }
    }
    st.sum[col] = S; st.comp[col] = C; st.cnt[col] = cnt; st.nz[col] = nz;
}

// Lists that were loaded for a key (resume) but did not advance while the key was silent are out of
// step with the lag's global head.  Before such a key takes part in the lock-step pushes again its
// column is rotated so that the list ends just before `head` (three strided reversals, one thread;
// this happens at most once per silent period of a key).
__device__ void rotate_column(double* col, size_t pitch, uint32_t L, uint32_t d) {
    auto rev = [&](uint32_t lo, uint32_t hi) {  // reverse [lo, hi)
        while (lo + 1 < hi) {
            --hi;
            const double a = col[(size_t)lo * pitch], b = col[(size_t)hi * pitch];
            col[(size_t)lo * pitch] = b;
            col[(size_t)hi * pitch] = a;
            ++lo;
        }
    };
    rev(0, L);
    rev(0, d);
    rev(d, L);  // new[(s + d) mod L] = old[s]
}

struct RingAlign {
    int num_lags;
    double* ring[APM_MAX_LAGS];
    const uint32_t* len[APM_MAX_LAGS];
    uint32_t* next[APM_MAX_LAGS];
    uint32_t L[APM_MAX_LAGS], head[APM_MAX_LAGS];
    size_t pitch;
};


// One thread per key, on the baseline stream right before the lag kernels of the same tick (so it
// is ordered against the pushes of the previous tick).
__global__ void __launch_bounds__(128) realign_kernel(const RingAlign ra, const uint8_t* __restrict__ emit, uint32_t K,
                                                      uint32_t* __restrict__ pending) {
    const uint32_t key = blockIdx.x * blockDim.x + threadIdx.x;
    if (key >= K) return;
    if (!emit[key]) {  // still silent: remember that it holds a list, so the host keeps this kernel in the tick
        bool holds = false;
        for (int l = 0; l < ra.num_lags; ++l)
            for (int s = 0; s < APM_NUM_SERIES; ++s) holds |= ra.len[l][(size_t)s * K + key] != 0;
        if (holds) atomicAdd(pending, 1u);
        return;
    }
    for (int l = 0; l < ra.num_lags; ++l) {
#pragma unroll 1
        for (int s = 0; s < APM_NUM_SERIES; ++s) {
            const size_t col = (size_t)s * K + key;
            if (ra.len[l][col] != 0 && ra.next[l][col] != ra.head[l]) {
                const uint32_t d = (ra.head[l] + ra.L[l] - ra.next[l][col]) % ra.L[l];
                rotate_column(ra.ring[l] + col, ra.pitch, ra.L[l], d);
                ra.next[l][col] = ra.head[l];
            }
        }
    }
}

// ---- alert counter core --------------------------------------------------------------------------
// processFSEntry (stream_process_alerts.js:348-434) for one (lag, key) per thread.  Its inputs are
// the fields an `fs` line carries after the toFixed hop, which the record already holds quantised.
struct AlertParams {
    apm_fs_record* fs;      // [num_lags][K] records of this tick
    uint32_t* count;        // [num_lags][K]
    const double* hard_max; // [K]
    const uint8_t* suppressed;  // [K]
    uint32_t K, num_lags;
    apm_alert_rules r;
};

__global__ void __launch_bounds__(256) alert_counter_kernel(const AlertParams p) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.K * p.num_lags) return;
    apm_fs_record* en = p.fs + i;
    if (!(en->flags & 1)) return;
    const uint32_t lag_idx = i / p.K, key = i - lag_idx * p.K;
    uint32_t cnt = p.count[i];
    bool incremented = false, trigger = false;
    uint32_t causes = 0;
    auto alert = [&](uint32_t cause) {
        if (!incremented) {  // only once per entry, and only while the rolling total is <= the window (:372-377)
            if (cnt <= (uint32_t)p.r.window) cnt++;
            incremented = true;
        }
        if (p.r.window > 1 && p.r.required > 1) {
            if (cnt >= (uint32_t)p.r.required) { trigger = true; causes |= cause; }
        } else {
            trigger = true;
            causes |= cause;
        }
    };
    if (!((p.r.suppressed_lag_mask >> lag_idx) & 1u) && !p.suppressed[key]) {
        const double hm = p.hard_max[key] > 0.0 ? p.hard_max[key] : p.r.hard_max_ms;
        const double avg = en->val[0], p75 = en->val[1], tpm = en->tpm;
        if (avg > hm) alert(APM_CAUSE_AVG_HARD_MAX);
        if (p75 > hm) alert(APM_CAUSE_P75_HARD_MAX);
        int both = 0;
        if (en->signal[0] > 0 && avg > p.r.hard_min_ms && tpm > p.r.hard_min_tpm) {
            if (!p.r.alert_on_both_only) alert(APM_CAUSE_AVG_UB); else both++;
        }
        if (en->signal[1] > 0 && p75 > p.r.hard_min_ms && tpm > p.r.hard_min_tpm) {
            if (!p.r.alert_on_both_only) alert(APM_CAUSE_P75_UB); else both++;
        }
        if (p.r.alert_on_both_only && both >= 2) alert(APM_CAUSE_BOTH_UB);
    }
    if (!incremented && cnt > 0) cnt--;  // (:427-429)
    p.count[i] = cnt;
    en->flags = (uint8_t)(1u | (trigger ? 2u : 0u));
    en->pad_[0] = (uint8_t)causes;
    en->pad_[1] = (uint8_t)(cnt > 255u ? 255u : cnt);
}

// Copies the fs records that carry a signal into the tick's compact alert block, in the order of the fs
// block itself ((lag, key)-ascending), so the block needs no sorting afterwards: every CTA counts the hits among
// its 256 records (pass 1), then adds up the counts of the CTAs before it, ranks its hits with a ballot and copies
// them (pass 2).  No counter has to be zeroed beforehand; the last CTA stores the total.
constexpr int CA_THREADS = 256;

__device__ __forceinline__ bool ca_hit(const apm_fs_record* fs, uint32_t i, uint32_t n, int by_trigger) {
    if (i >= n) return false;
    const apm_fs_record* r = fs + i;
    if (!(r->flags & 1)) return false;
    return by_trigger ? (r->flags & 2) != 0 : (r->signal[0] | r->signal[1] | r->signal[2]) != 0;
}

__global__ void __launch_bounds__(CA_THREADS) count_alerts_kernel(const apm_fs_record* __restrict__ fs, uint32_t n, int by_trigger,
                                                                  uint32_t* __restrict__ block_counts) {
    const uint32_t c = __syncthreads_count(ca_hit(fs, blockIdx.x * CA_THREADS + threadIdx.x, n, by_trigger));
    if (threadIdx.x == 0) block_counts[blockIdx.x] = c;
}

__global__ void __launch_bounds__(CA_THREADS) compact_alerts_kernel(const apm_fs_record* __restrict__ fs, uint32_t n,
                                                                    apm_fs_record* __restrict__ alerts,
                                                                    unsigned int* __restrict__ count, uint32_t cap, int by_trigger,
                                                                    const uint32_t* __restrict__ block_counts) {
    __shared__ uint32_t s_warp[CA_THREADS / 32];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // hits in the CTAs before this one
    uint32_t part = 0;
    for (uint32_t j = tid; j < blockIdx.x; j += CA_THREADS) part += block_counts[j];
    part = __reduce_add_sync(0xffffffffu, part);
    if (lane == 0) s_warp[w] = part;
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int j = 0; j < CA_THREADS / 32; ++j) t += s_warp[j];
        s_base = t;
    }
    __syncthreads();
    const uint32_t base = s_base;
    __syncthreads();
    const uint32_t i = blockIdx.x * CA_THREADS + tid;
    const bool hit = ca_hit(fs, i, n, by_trigger);
    const uint32_t bal = __ballot_sync(0xffffffffu, hit);
    if (lane == 0) s_warp[w] = __popc(bal);
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int j = 0; j < CA_THREADS / 32; ++j) {
        const uint32_t c = s_warp[j];
        if (j < w) before += c;
        total += c;
    }
    const uint32_t k = base + before + __popc(bal & ((1u << lane) - 1u));
    if (hit && k < cap) {
        const uint4* src = reinterpret_cast<const uint4*>(fs + i);
        uint4* dst = reinterpret_cast<uint4*>(alerts + k);
#pragma unroll
        for (int j = 0; j < 8; ++j) dst[j] = src[j];
    }
    if (blockIdx.x == gridDim.x - 1 && tid == 0) *count = base + total;
}

// The same compaction for a few thousand records (the reference's deployment: 400 keys x 2 lags) in one launch: a
// single CTA walks the records 1024 at a time with a running base, in record order.
__global__ void __launch_bounds__(1024) compact_alerts_small_kernel(const apm_fs_record* __restrict__ fs, uint32_t n,
                                                                    apm_fs_record* __restrict__ alerts, unsigned int* __restrict__ count,
                                                                    uint32_t cap, int by_trigger) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (uint32_t i0 = 0; i0 < n; i0 += 1024) {
        const uint32_t i = i0 + tid;
        const bool hit = ca_hit(fs, i, n, by_trigger);
        const uint32_t bal = __ballot_sync(0xffffffffu, hit);
        if (lane == 0) s_warp[w] = __popc(bal);
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const uint32_t c = s_warp[j];
            if (j < w) before += c;
            total += c;
        }
        const uint32_t k = s_base + before + __popc(bal & ((1u << lane) - 1u));
        if (hit && k < cap) {
            const uint4* src = reinterpret_cast<const uint4*>(fs + i);
            uint4* dst = reinterpret_cast<uint4*>(alerts + k);
#pragma unroll
            for (int j = 0; j < 8; ++j) dst[j] = src[j];
        }
        __syncthreads();
        if (tid == 0) s_base += total;
        __syncthreads();
    }
    if (tid == 0) *count = s_base;
}

}  // namespace

apm_status apm_baseline_make_tmaps(apm_ctx* c) {
    c->ring_tmap_ok = false;
    typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn ||
        qres != cudaDriverEntryPointSuccess)
        return apm_fail(c, APM_E_DEVICE, "cuTensorMapEncodeTiled is not available in this driver");
    for (int l = 0; l < c->cfg.num_lags; ++l) {
        const cuuint64_t gdim[2] = {(cuuint64_t)c->ring_pitch, (cuuint64_t)c->cfg.lags[l]};
        const cuuint64_t gstride[1] = {(cuuint64_t)(c->ring_pitch * sizeof(double))};
        const cuuint32_t box[2] = {K4T_COLS, K4T_ROWS};
        const cuuint32_t estr[2] = {1, 1};
        CUresult r = ((EncodeTiled)fn)(&c->ring_tmap[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, c->ring[l], gdim, gstride, box, estr,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return apm_fail(c, APM_E_DEVICE, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    }
    CUDA_TRY(c, cudaFuncSetAttribute(baseline_tma_kernel<K4T_DEEP_STAGES, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(K4TSmem<K4T_DEEP_STAGES, 3>) + 128));
    for (int l = 0; l < c->cfg.num_lags; ++l) {  // the few-list walk loads 96 slots per request
        const cuuint64_t gdim[2] = {(cuuint64_t)c->ring_pitch, (cuuint64_t)c->cfg.lags[l]};
        const cuuint64_t gstride[1] = {(cuuint64_t)(c->ring_pitch * sizeof(double))};
        const cuuint32_t box[2] = {K4T_COLS, K4F_ROWS};
        const cuuint32_t estr[2] = {1, 1};
        CUresult r = ((EncodeTiled)fn)(&c->ring_tmap_few[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, c->ring[l], gdim, gstride, box, estr,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return apm_fail(c, APM_E_DEVICE, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    }
    CUDA_TRY(c, cudaFuncSetAttribute(baseline_few_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K4FSmem) + 128));
    c->ring_tmap_ok = true;
    return APM_OK;
}

// Few lists: every lag's column groups together fit the SMs one CTA each (cfg1: 2 x 38 of 148) - or twice that: the
// CTAs do not wait for each other, the longest lag's are placed first, and two waves of the latency-bound walk (2 x 50 us
// for day-long lists) still beat the bandwidth-oriented kernel, whose one warp per CTA takes 113 us for such a list.
bool apm_baseline_few_ok(const apm_ctx* c) {
    if (!c->k4_few || !c->ring_tmap_ok || c->cfg.sd_mode != APM_SD_REFERENCE || c->k4_force_ldg || c->k4_stages != 0) return false;
    if (c->cfg.flags & APM_F_INCREMENTAL) return false;
    const unsigned g = (3u * c->K + K4T_COLS - 1) / K4T_COLS;
    return c->cfg.num_lags > 0 && g * (unsigned)c->cfg.num_lags <= (unsigned)c->k4_few_waves * (unsigned)c->num_sms;
}

// All lags of one tick in one launch (see baseline_few_kernel).
apm_status apm_launch_baseline_few(apm_ctx* c, int64_t tick_ts, uint32_t tick_slot, int buf, int alert_slot, cudaStream_t st) {
    ProfScope prof_scope_(c, APM_PG_BASELINE, st);
    FewParams fp;
    memset(&fp, 0, sizeof(fp));
    const unsigned g = (3u * c->K + K4T_COLS - 1) / K4T_COLS;
    fp.n = c->cfg.num_lags;
    for (int l = 0; l < c->cfg.num_lags; ++l) {
        BaselineParams& p = fp.p[l];
        fp.tmap[l] = c->ring_tmap_few[l];
        p.ring = c->ring[l]; p.ring_len = c->ring_len[l]; p.ring_next = c->ring_next[l]; p.zin = c->zin[buf]; p.thr = c->thr[l]; p.infl = c->infl[l];
        p.seen = c->emit[buf];
        p.fs = c->d_fs ? c->d_fs + ((size_t)tick_slot * c->cfg.num_lags + l) * c->K : nullptr;
        p.K = c->K; p.L = (uint32_t)c->cfg.lags[l]; p.head = c->ring_head[l]; p.lag_value = (uint32_t)c->cfg.lags[l];
        p.pitch = (uint32_t)c->ring_pitch;
        p.tick_ts = tick_ts;
        p.sd_mode = c->cfg.sd_mode;
        // lags are ascending (apm_config): launch order is descending, so the longest walk's CTAs are placed first
        fp.lag_of[c->cfg.num_lags - 1 - l] = l;
        fp.cta_end[l] = g * (unsigned)(l + 1);
    }
    if (alert_slot >= 0) {
        fp.compact_n = c->K * (uint32_t)c->cfg.num_lags;
        fp.alert_cap = c->cfg.max_alerts_per_tick;
        fp.fs_tick = c->d_fs + (size_t)tick_slot * c->cfg.num_lags * c->K;
        fp.alerts = c->d_alerts + (size_t)alert_slot * c->cfg.max_alerts_per_tick;
        fp.alert_count = c->d_alert_count + alert_slot;
        fp.done = c->k4_done;
    }
    baseline_few_kernel<<<g * (unsigned)c->cfg.num_lags, 96, sizeof(K4FSmem) + 128, st>>>(fp);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    for (int l = 0; l < c->cfg.num_lags; ++l) c->ring_head[l] = (c->ring_head[l] + 1) % (uint32_t)c->cfg.lags[l];
    return APM_OK;
}

apm_status apm_launch_baseline(apm_ctx* c, int l, int64_t tick_ts, uint32_t tick_slot, int buf, cudaStream_t st) {
    ProfScope prof_scope_(c, APM_PG_BASELINE, st);
    BaselineParams p;
    p.ring = c->ring[l]; p.ring_len = c->ring_len[l]; p.ring_next = c->ring_next[l]; p.zin = c->zin[buf]; p.thr = c->thr[l]; p.infl = c->infl[l];
    p.seen = c->emit[buf];
    p.fs = c->d_fs ? c->d_fs + ((size_t)tick_slot * c->cfg.num_lags + l) * c->K : nullptr;
    p.K = c->K; p.L = (uint32_t)c->cfg.lags[l]; p.head = c->ring_head[l]; p.lag_value = (uint32_t)c->cfg.lags[l];
    p.pitch = (uint32_t)c->ring_pitch;
    p.tick_ts = tick_ts;
    p.sd_mode = c->cfg.sd_mode;
    const uint32_t ncol = 3u * c->K;
    if ((c->cfg.flags & APM_F_INCREMENTAL) && c->cfg.sd_mode == APM_SD_REFERENCE) {
        IncState is = {c->inc_sum[l], c->inc_comp[l], c->inc_cnt[l], c->inc_nz[l]};
        if (c->inc_dirty[l]) {  // lists were loaded since the last tick: rebuild the running statistics first
            inc_init_kernel<<<(ncol + K4_THREADS - 1) / K4_THREADS, K4_THREADS, 0, st>>>(c->ring[l], c->ring_len[l], c->ring_next[l], is,
                                                                                   c->K, p.L, c->ring_pitch);
            c->inc_dirty[l] = false;
            c->launches += 1;
        }
        baseline_inc_kernel<<<(ncol + K4_THREADS - 1) / K4_THREADS, K4_THREADS, 0, st>>>(p, is, &c->d_counters->inc_fallbacks);
    } else if (c->ring_tmap_ok && c->cfg.sd_mode == APM_SD_REFERENCE && !c->k4_force_ldg) {
        const unsigned g = (ncol + K4T_COLS - 1) / K4T_COLS;
        // With few column groups (a few hundred keys) one warp per SM cannot keep enough bytes in
        // flight with 4 x 4 KB stages: the list walk is then bound by load latency, not by HBM.
        // Below two CTAs per SM the deep instance (24 stages, 96 KB in flight per CTA) is used.
        int stages = c->k4_stages;
        if (stages == 0) stages = g <= 2u * (unsigned)c->num_sms ? K4T_DEEP_STAGES : 4;
#define APM_K4T_LAUNCH(NS, M) baseline_tma_kernel<NS, M><<<g, 32, sizeof(K4TSmem<NS, M>) + 128, st>>>(c->ring_tmap[l], p)
        switch (stages) {
            case 2: APM_K4T_LAUNCH(2, 1); break;
            case 3: APM_K4T_LAUNCH(3, 1); break;
            case 6: APM_K4T_LAUNCH(6, 1); break;
            case K4T_DEEP_STAGES: APM_K4T_LAUNCH(K4T_DEEP_STAGES, 3); break;
            default: APM_K4T_LAUNCH(4, 1); break;
        }
#undef APM_K4T_LAUNCH
    } else {
        baseline_kernel<<<(ncol + K4_THREADS - 1) / K4_THREADS, K4_THREADS, 0, st>>>(p);
    }
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    c->ring_head[l] = (c->ring_head[l] + 1) % p.L;
    return APM_OK;
}

apm_status apm_launch_realign(apm_ctx* c, int buf, cudaStream_t st) {
    ProfScope prof_scope_(c, APM_PG_REALIGN, st);
    RingAlign ra;
    ra.num_lags = c->cfg.num_lags;
    ra.pitch = c->ring_pitch;
    for (int l = 0; l < ra.num_lags; ++l) {
        ra.ring[l] = c->ring[l]; ra.len[l] = c->ring_len[l]; ra.next[l] = c->ring_next[l];
        ra.L[l] = (uint32_t)c->cfg.lags[l]; ra.head[l] = c->ring_head[l];
    }
    CUDA_TRY(c, cudaMemsetAsync(&c->d_counters->pending_lists, 0, sizeof(uint32_t), st));
    realign_kernel<<<(c->K + 127) / 128, 128, 0, st>>>(ra, c->emit[buf], c->K, &c->d_counters->pending_lists);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

apm_status apm_launch_alert_counter(apm_ctx* c, uint32_t fs_slot, cudaStream_t st) {
    if (!c->alert_rules.enabled || !c->d_fs) return APM_OK;
    ProfScope prof_scope_(c, APM_PG_ALERTS, st);
    AlertParams p;
    p.fs = c->d_fs + (size_t)fs_slot * c->cfg.num_lags * c->K;
    p.count = c->alert_count; p.hard_max = c->alert_hard_max; p.suppressed = c->alert_suppressed;
    p.K = c->K; p.num_lags = (uint32_t)c->cfg.num_lags; p.r = c->alert_rules;
    const uint32_t n = c->K * (uint32_t)c->cfg.num_lags;
    alert_counter_kernel<<<(n + 255) / 256, 256, 0, st>>>(p);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

apm_status apm_launch_compact_alerts(apm_ctx* c, uint32_t fs_slot, uint32_t alert_slot, cudaStream_t st) {
    ProfScope prof_scope_(c, APM_PG_ALERTS, st);
    if (!c->d_alerts || !c->d_fs) return APM_OK;
    const uint32_t n = c->K * (uint32_t)c->cfg.num_lags;
    const unsigned g = (n + CA_THREADS - 1) / CA_THREADS;
    const apm_fs_record* fs = c->d_fs + (size_t)fs_slot * c->cfg.num_lags * c->K;
    if (n <= 4096 && c->small_tick) {
        compact_alerts_small_kernel<<<1, 1024, 0, st>>>(fs, n, c->d_alerts + (size_t)alert_slot * c->cfg.max_alerts_per_tick,
                                                        c->d_alert_count + alert_slot, c->cfg.max_alerts_per_tick, c->alert_rules.enabled);
        c->launches += 1;
        CUDA_TRY(c, cudaGetLastError());
        return APM_OK;
    }
    count_alerts_kernel<<<g, CA_THREADS, 0, st>>>(fs, n, c->alert_rules.enabled, c->d_alert_blocks);
    compact_alerts_kernel<<<g, CA_THREADS, 0, st>>>(fs, n, c->d_alerts + (size_t)alert_slot * c->cfg.max_alerts_per_tick,
                                                    c->d_alert_count + alert_slot, c->cfg.max_alerts_per_tick, c->alert_rules.enabled,
                                                    c->d_alert_blocks);
    c->launches += 2;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k4_baseline() {
    const void* fns[] = {(const void*)alert_counter_kernel, (const void*)baseline_inc_kernel, (const void*)baseline_kernel, (const void*)baseline_tma_kernel<2, 1>, (const void*)baseline_tma_kernel<3, 1>, (const void*)baseline_tma_kernel<4, 1>, (const void*)baseline_tma_kernel<6, 1>, (const void*)baseline_tma_kernel<K4T_DEEP_STAGES, 3>, (const void*)baseline_few_kernel, (const void*)compact_alerts_kernel, (const void*)compact_alerts_small_kernel, (const void*)count_alerts_kernel, (const void*)inc_init_kernel, (const void*)realign_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}

#ifdef APM_K4_DEBUG
extern "C" int apm_debug_k4(unsigned long long* out32) {
    return (int)cudaMemcpyFromSymbol(out32, apm_k4_dbg, sizeof(apm_k4_dbg));
}
#endif
