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
#include "ctx.h"

namespace {

constexpr int K4_THREADS = 128;
constexpr int K4_UNROLL = 32;

struct BaselineParams {
    double* ring;          // [L][3K]
    uint32_t* ring_len;    // [3K]
    const double* zin;     // [4][K]: tpm', val0..2
    const double* thr;     // [K]
    const double* infl;    // [K]
    const uint8_t* seen;   // [K]
    apm_fs_record* fs;     // [K] records of this (tick, lag); nullptr when not retained
    uint32_t K, L, head, lag_value;
    int64_t tick_ts;
    int sd_mode;
};

__device__ __forceinline__ void accumulate(double v, double& sum, uint32_t& cnt) {
    // average(): `value != null && !isNaN(parseFloat(value))` then `sum + parseFloat(value)`
    if (v == v) {
        sum = __dadd_rn(sum, v);
        cnt += 1;
    }
}

__global__ void __launch_bounds__(K4_THREADS) baseline_kernel(const BaselineParams p) {
    const uint32_t col = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t ncol = 3u * p.K;
    if (col >= ncol) return;
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

// Copies the fs records that carry a signal into the tick's compact alert block.
__global__ void __launch_bounds__(256) compact_alerts_kernel(const apm_fs_record* __restrict__ fs, uint32_t n,
                                                             apm_fs_record* __restrict__ alerts,
                                                             unsigned int* __restrict__ count, uint32_t cap) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const apm_fs_record* r = fs + i;
    if (!(r->flags & 1)) return;
    if (r->signal[0] | r->signal[1] | r->signal[2]) {
        const unsigned int k = atomicAdd(count, 1u);
        if (k < cap) {
            const uint4* s = reinterpret_cast<const uint4*>(r);
            uint4* d = reinterpret_cast<uint4*>(alerts + k);
#pragma unroll
            for (int j = 0; j < 8; ++j) d[j] = s[j];
        }
    }
}

}  // namespace

apm_status apm_launch_baseline(apm_ctx* c, int l, int64_t tick_ts, uint32_t tick_slot) {
    ProfScope prof_scope_(c, APM_PG_BASELINE);
    BaselineParams p;
    p.ring = c->ring[l]; p.ring_len = c->ring_len[l]; p.zin = c->zin; p.thr = c->thr[l]; p.infl = c->infl[l];
    p.seen = c->seen;
    p.fs = c->d_fs ? c->d_fs + ((size_t)tick_slot * c->cfg.num_lags + l) * c->K : nullptr;
    p.K = c->K; p.L = (uint32_t)c->cfg.lags[l]; p.head = c->ring_head[l]; p.lag_value = (uint32_t)c->cfg.lags[l];
    p.tick_ts = tick_ts;
    p.sd_mode = c->cfg.sd_mode;
    const uint32_t ncol = 3u * c->K;
    baseline_kernel<<<(ncol + K4_THREADS - 1) / K4_THREADS, K4_THREADS, 0, c->s_main>>>(p);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    c->ring_head[l] = (c->ring_head[l] + 1) % p.L;
    return APM_OK;
}

apm_status apm_launch_compact_alerts(apm_ctx* c, uint32_t fs_slot, uint32_t alert_slot) {
    ProfScope prof_scope_(c, APM_PG_ALERTS);
    if (!c->d_alerts || !c->d_fs) return APM_OK;
    const uint32_t n = c->K * (uint32_t)c->cfg.num_lags;
    compact_alerts_kernel<<<(n + 255) / 256, 256, 0, c->s_main>>>(
        c->d_fs + (size_t)fs_slot * c->cfg.num_lags * c->K, n,
        c->d_alerts + (size_t)alert_slot * c->cfg.max_alerts_per_tick, c->d_alert_count + alert_slot,
        c->cfg.max_alerts_per_tick);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}
