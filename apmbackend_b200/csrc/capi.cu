// extern "C" surface of libapm_b200.so (include/apm_b200.h) and the host-side tick logic.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "ctx.h"

static thread_local std::string g_create_error;

apm_status apm_fail(apm_ctx* ctx, apm_status st, const std::string& msg) {
    if (ctx) ctx->err = msg; else g_create_error = msg;
    return st;
}

apm_status apm_fail_cuda(apm_ctx* ctx, cudaError_t e, const char* what, const char* file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    if (ctx) { ctx->err = buf; ctx->device_error = true; } else g_create_error = buf;
    return APM_E_DEVICE;
}

int apm_sm_count() {
    static int cached[APM_MAX_DEVICES];  // 0 = not asked yet
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= APM_MAX_DEVICES) return 148;
    int n = cached[dev];
    if (n <= 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return n;
}

bool apm_first_on_device(unsigned long long* done) {
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= APM_MAX_DEVICES) return true;
    std::lock_guard<std::mutex> lock(mu);
    if (*done & (1ull << dev)) return false;
    *done |= 1ull << dev;
    return true;
}

namespace {

template <typename T>
cudaError_t dalloc(T** p, size_t n) {
    *p = nullptr;
    if (n == 0) return cudaSuccess;
    return cudaMalloc((void**)p, n * sizeof(T));
}

// 12-byte transactions (apm_tx_packed) -> the 20-byte columns every kernel reads
__global__ void __launch_bounds__(256) widen_tx_kernel(const uint32_t* __restrict__ key, const uint32_t* __restrict__ dt,
                                                       const uint32_t* __restrict__ el, int64_t base, uint64_t n, apm_tx_columns d) {
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const uint32_t k = __ldcs(key + i);
    d.service_id[i] = k & 0xFFFFu;
    d.jvm_id[i] = k >> 16;
    d.end_ts_ms[i] = base + (int64_t)__ldcs(dt + i);
    d.elapsed_ms[i] = (int32_t)__ldcs(el + i);
}

// 8-byte transactions (apm_process_host_packed8) -> the same columns.  A word whose elapsed field is all ones looks its
// duration up in the call's escape list (sorted by transaction index, a handful of entries at most).
__global__ void __launch_bounds__(256) widen_tx8_kernel(const uint64_t* __restrict__ tx, const apm_tx_escape* __restrict__ esc, uint32_t num_esc,
                                                        uint64_t first_index, int64_t base, uint64_t n, uint32_t num_services,
                                                        apm_tx_columns d) {
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const uint64_t w = __ldcs(tx + i);
    const uint32_t key = (uint32_t)(w >> 48), dt = (uint32_t)(w >> 24) & 0xFFFFFFu;
    uint32_t el = (uint32_t)w & 0xFFFFFFu;
    if (el == 0xFFFFFFu) {
        // not in the list: the field stands for itself (a duration of exactly 16 777 215 ms)
        const uint64_t idx = first_index + i;
        uint32_t lo = 0, hi = num_esc;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if ((uint64_t)esc[mid].index < idx) lo = mid + 1; else hi = mid;
        }
        if (lo < num_esc && (uint64_t)esc[lo].index == idx) el = (uint32_t)esc[lo].elapsed_ms;
    }
    // key -> (jvm, service): a key at or above num_services * num_jvms becomes a jvm id out of range, i.e. an invalid transaction
    d.service_id[i] = key % num_services;
    d.jvm_id[i] = key / num_services;
    d.end_ts_ms[i] = base + (int64_t)dt;
    d.elapsed_ms[i] = (int32_t)el;
}

apm_status alloc_columns_device(apm_ctx* c, apm_tx_columns* d, uint64_t n) {
    CUDA_TRY(c, dalloc(&d->service_id, n));
    CUDA_TRY(c, dalloc(&d->jvm_id, n));
    CUDA_TRY(c, dalloc(&d->end_ts_ms, n));
    CUDA_TRY(c, dalloc(&d->elapsed_ms, n));
    return APM_OK;
}

void free_columns_device(apm_tx_columns* d) {
    cudaFree(d->service_id); cudaFree(d->jvm_id); cudaFree(d->end_ts_ms); cudaFree(d->elapsed_ms);
    *d = {nullptr, nullptr, nullptr, nullptr};
}

apm_status validate(const apm_config* cfg) {
    if (!cfg) return apm_fail(nullptr, APM_E_INVALID, "cfg is NULL");
    if (cfg->struct_size != sizeof(apm_config)) return apm_fail(nullptr, APM_E_INVALID, "apm_config.struct_size mismatch");
    if (cfg->num_services == 0 || cfg->num_jvms == 0) return apm_fail(nullptr, APM_E_INVALID, "num_services/num_jvms must be > 0");
    if ((uint64_t)cfg->num_services * cfg->num_jvms > (1ull << 26)) return apm_fail(nullptr, APM_E_INVALID, "too many keys (limit 2^26 per GPU)");
    if (cfg->bucket_ms <= 0) return apm_fail(nullptr, APM_E_INVALID, "bucket_ms must be > 0");
    if (cfg->window_buckets < 1 || cfg->buffer_buckets < 0 || cfg->window_buckets + cfg->buffer_buckets + 1 > APM_MAX_SLOTS)
        return apm_fail(nullptr, APM_E_INVALID, "window/buffer out of range (window >= 1, window+buffer+1 <= 64)");
    if (cfg->num_percentiles < 0 || cfg->num_percentiles > APM_MAX_PCT) return apm_fail(nullptr, APM_E_INVALID, "num_percentiles out of range");
    for (int i = 0; i < cfg->num_percentiles; ++i)
        if (!(cfg->percentiles[i] >= 0.0 && cfg->percentiles[i] <= 100.0)) return apm_fail(nullptr, APM_E_INVALID, "percentile out of [0,100]");
    if (cfg->num_lags < 0 || cfg->num_lags > APM_MAX_LAGS) return apm_fail(nullptr, APM_E_INVALID, "num_lags out of range");
    for (int i = 0; i < cfg->num_lags; ++i) {
        if (cfg->lags[i] < 1) return apm_fail(nullptr, APM_E_INVALID, "lag must be >= 1");
        if (i && cfg->lags[i] <= cfg->lags[i - 1]) return apm_fail(nullptr, APM_E_INVALID, "lags must be ascending");
    }
    if (cfg->max_batch == 0) return apm_fail(nullptr, APM_E_INVALID, "max_batch must be > 0");
    if (cfg->out_ticks == 0) return apm_fail(nullptr, APM_E_INVALID, "out_ticks must be > 0");
    if ((cfg->flags & APM_F_PERCENTILES) && cfg->num_percentiles > 0 && cfg->max_tx_per_bucket == 0)
        return apm_fail(nullptr, APM_E_INVALID, "max_tx_per_bucket must be > 0 with percentiles");
    return APM_OK;
}

// waits for everything queued on the context (ingest/sort stream, baseline stream, copies)
apm_status sync_all(apm_ctx* c) {
    apm_status fl = apm_flush_pending(c);
    if (fl) return fl;
    CUDA_TRY(c, cudaStreamSynchronize(c->s_copy));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    return APM_OK;
}

bool with_pct(const apm_ctx* c) { return (c->cfg.flags & APM_F_PERCENTILES) && c->cfg.num_percentiles > 0; }
bool with_z(const apm_ctx* c) { return (c->cfg.flags & APM_F_ZSCORE) && c->cfg.num_lags > 0; }

apm_status refresh_counters(apm_ctx* c) {
    apm_status fl = apm_flush_pending(c);
    if (fl) return fl;
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));  // first: the baseline stream also writes a counter (pending_lists)
    CUDA_TRY(c, cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    c->counters.tx_ingested = c->h_counters->ingested;
    c->counters.tx_dropped = c->h_counters->dropped;
    c->counters.tx_invalid = c->h_counters->invalid + c->h_counters->future;
    c->counters.kernel_launches = c->launches;
    c->counters.inc_fallbacks = c->h_counters->inc_fallbacks;
    if (c->may_misalign && c->tick_seq > c->misalign_mark && c->h_counters->pending_lists == 0) c->may_misalign = false;
    if (c->h_counters->log_overflow)
        return apm_fail(c, APM_E_CAPACITY, "a device buffer sized by max_tx_per_bucket overflowed (a bucket's value log or the passthrough reorder buffer); results that depend on it are incomplete - recreate the context with a larger capacity");
    return APM_OK;
}

// The z-score stage of one tick: realignment (when lists were loaded), one kernel per lag, alert compaction.
apm_status launch_baseline_group(apm_ctx* c, const PendingBaseline& pb, cudaStream_t sb) {
    apm_status st;
    if (c->may_misalign) {
        st = apm_launch_realign(c, pb.buf, sb);  // lists of keys that were silent re-enter the lock step
        if (st) return st;
    }
    bool compacted = false;
    if (apm_baseline_few_ok(c)) {  // a few hundred lists: every lag in one launch, loader and adder warps
        // ... and, when the signalled records themselves are the alerts (no stage-[4] rules), the last CTA to finish
        // compacts them: one launch less on the baseline stream, which is the tick's critical path at this shape
        compacted = (c->cfg.flags & APM_F_ALERTS) && !c->alert_rules.enabled && c->d_alerts && c->d_fs && c->k4_done && c->small_tick &&
                    (uint64_t)c->K * c->cfg.num_lags <= 4096;
        st = apm_launch_baseline_few(c, pb.tick_ts, pb.fslot, pb.buf, compacted ? (int)pb.tslot : -1, sb);
        if (st) return st;
    } else {
        for (int l = 0; l < c->cfg.num_lags; ++l) {  // Object.entries(lags): ascending (:217)
            st = apm_launch_baseline(c, l, pb.tick_ts, pb.fslot, pb.buf, sb);
            if (st) return st;
        }
    }
    st = apm_launch_alert_counter(c, pb.fslot, sb);  // processFSEntry's counter, when rules are set
    if (st) return st;
    if (c->cfg.flags & APM_F_ALERTS) {
        if (!compacted) {
            st = apm_launch_compact_alerts(c, pb.fslot, pb.tslot, sb);
            if (st) return st;
        }
        if (c->h_gather_bound) {  // all-gathered while the next ticks run
            st = apm_comm_tick_gather(c, pb.tslot, sb);
            if (st) return st;
        }
    }
    if (c->h_fs_bound && (c->cfg.flags & APM_F_KEEP_FS)) {
        // this tick's records go to the bound pinned buffer on their own stream, under the following ticks
        const size_t per_tick = (size_t)c->K * c->cfg.num_lags;
        if (((size_t)pb.tslot + 1) * per_tick <= c->h_fs_bound_cap) {
            CUDA_TRY(c, cudaEventRecord(c->ev_d2h, sb));
            CUDA_TRY(c, cudaStreamWaitEvent(c->s_d2h, c->ev_d2h, 0));
            CUDA_TRY(c, cudaMemcpyAsync(c->h_fs_bound + (size_t)pb.tslot * per_tick, c->d_fs + (size_t)pb.fslot * per_tick,
                                        per_tick * sizeof(apm_fs_record), cudaMemcpyDeviceToHost, c->s_d2h));
            c->counters.d2h_bytes += per_tick * sizeof(apm_fs_record);
        }
    }
    return APM_OK;
}

// host-located ticks are consumed by one process_* call, whether it succeeds or not
struct ExtMarksScope {
    apm_ctx* c;
    ~ExtMarksScope() { c->ext_armed = false; c->ext_marks.clear(); c->ext_next = 0; }
};

// The tick itself: everything consumeMsg does inside `if (bucketLabel > latestBucket)`
// (stream_calc_stats.js:348-367) and, behind it, processData for each emitted StatEntry.
apm_status do_tick(apm_ctx* c, int64_t new_latest) {
    if (new_latest <= c->latest) return apm_fail(c, APM_E_INVALID, "tick label must exceed latestBucket");
    apm_status fl = apm_flush_pending(c);  // the previous tick's baseline goes in behind whatever was ingested since
    if (fl) return fl;
    const bool keep_any = c->cfg.flags & (APM_F_KEEP_STATS | APM_F_KEEP_FS | APM_F_ALERTS);
    if (keep_any && c->ticks_buffered >= c->cfg.out_ticks) return apm_fail(c, APM_E_WOULDBLOCK, "output ring full: fetch results first");
    const int64_t old_latest = c->latest;
    c->latest = new_latest;  // :351
    // removeOldBuckets (:103-113): labels below latest - keep are dropped; their slots take the new labels
    uint64_t mask = 0;
    const int64_t first_new = std::max(old_latest + 1, new_latest - c->S + 1);
    for (int64_t lab = first_new; lab <= new_latest; ++lab) {
        if (lab < 0) continue;
        const int slot = (int)(lab % c->S);
        mask |= 1ull << slot;
        c->slot_label[slot] = lab;
        c->host_count[slot] = c->host_sorted[slot] = 0;
    }
    // window = labels in [latest - keep, latest - buffer] (:172)
    WindowSlots w;
    w.n = 0;
    for (int64_t lab = new_latest - c->keep; lab <= new_latest - c->cfg.buffer_buckets; ++lab) {
        if (lab < 0) continue;
        const int slot = (int)(lab % c->S);
        if (c->slot_label[slot] != lab) continue;  // never filled
        w.slot[w.n++] = slot;
    }
    // the window's value logs that grew since their last sort; their lengths come from the chunk's segment table
    // (the chunk's pre-pass), not from the device, so nothing here waits for the GPU
    int grown[APM_MAX_SLOTS], ngrown = 0;
    uint64_t grown_n[APM_MAX_SLOTS], largest = 0;
    if (with_pct(c)) {
        for (int i = 0; i < w.n; ++i) {
            const int slot = w.slot[i];
            const uint64_t count = c->host_count[slot];
            if (count > c->log_cap) return apm_fail(c, APM_E_CAPACITY, "a bucket's value log overflowed max_tx_per_bucket");
            if (count != c->host_sorted[slot]) {
                grown[ngrown] = slot; grown_n[ngrown] = count; ++ngrown;
                largest = std::max(largest, count);
            }
        }
    }
    apm_status st;
    if (with_pct(c) && !c->pt_cap && apm_small_tick_ok(c, largest)) {
        // small shapes (the reference's own deployment): slot clear, sorts and accumulator rebuilds are one launch
        st = apm_small_tick(c, grown, grown_n, ngrown, mask);
        if (st) return st;
        for (int i = 0; i < ngrown; ++i) c->host_sorted[grown[i]] = grown_n[i];
    } else {
        st = apm_launch_clear_slots(c, mask);
        if (st) return st;
        if (c->pt_cap) {  // writeHeapToQueue(edgeTs) comes before generateAllStatsToQueue (:364-366)
            if (c->pt_base_ms == INT64_MIN) c->pt_base_ms = new_latest * c->cfg.bucket_ms - (1ll << 39);
            st = apm_pt_tick(c, (new_latest - c->cfg.buffer_buckets - 1) * c->cfg.bucket_ms, keep_any ? c->ticks_buffered : 0);
            if (st) return st;
        }
        for (int i = 0; i < ngrown; ++i) {
            st = apm_sort_slot_n(c, grown[i], grown_n[i], c->known_max_value);
            if (st) return st;
            st = apm_acc_from_sorted(c, grown[i], grown_n[i]);  // count / sum / sum of squares / min / max from the runs
            if (st) return st;
            c->host_sorted[grown[i]] = grown_n[i];
        }
    }
    const int64_t tick_ts = (new_latest - c->cfg.buffer_buckets - 1) * c->cfg.bucket_ms;  // edgeTs (:356)
    const uint32_t tslot = keep_any ? c->ticks_buffered : 0;
    const uint64_t seq = c->tick_seq++;
    const int buf = (int)(seq & 1), ev = (int)(seq & 3);
    // the window kernel overwrites zin/emit[buf], last read by the baseline of two ticks ago
    if (c->overlap && with_z(c) && seq >= 2) CUDA_TRY(c, cudaStreamWaitEvent(c->s_main, c->ev_base[(seq - 2) & 3], 0));
    st = apm_launch_tick_stats(c, w, tick_ts, (c->cfg.flags & APM_F_KEEP_STATS) ? tslot : 0, buf);
    if (st) return st;
    if (with_z(c)) {
        PendingBaseline pb = {true, tick_ts, (c->cfg.flags & APM_F_KEEP_FS) ? tslot : 0, tslot, buf, ev};
        if (c->overlap) {
            // Launched later - behind the next segment's ingest kernel (see apm_flush_pending) - so that the
            // bandwidth-hungry baseline does not start on top of the latency-bound head of the next tick's chain.
            // (no event here: apm_flush_pending records one on this stream behind the NEXT segment's ingest kernel, which
            // follows this tick's window kernel)
            c->pending = pb;
        } else {
            st = launch_baseline_group(c, pb, c->s_main);
            if (st) return st;
        }
    }
    if (keep_any) c->ticks_buffered++;
    c->counters.ticks++;
    return APM_OK;
}

// What segment i of a chunk appends to the value logs, by bucket age under the latestBucket in force
// (c->latest): keeps host_count equal to the device's SlotMeta.count.
void apply_seg_counts(apm_ctx* c, const apm_ctx::PrePass& pp, unsigned int i) {
    const uint32_t* row = pp.h_seg_counts + (size_t)i * (c->keep + 1);
    for (int age = 0; age <= c->keep; ++age)
        if (row[age]) c->host_count[(c->latest - age) % c->S] += row[age];
}

// latestBucket after a chunk whose pre-pass has finished: the label of its last tick
int64_t latest_after(const apm_ctx::PrePass& pp) { return pp.nt ? pp.h_ticks[pp.nt - 1].label : pp.latest0; }

// The ticks of one chunk, queued without waiting for the device: everything the loop needs from the data (where the
// ticks are, what each segment appends to each bucket) is in the chunk's finished pre-pass.
apm_status run_chunk(apm_ctx* c, const apm_tx_columns& d, uint64_t n, const apm_ctx::PrePass& pp, apm_tick_result* res) {
    const unsigned int nt = pp.nt;
    if (pp.latest0 != c->latest) return apm_fail(c, APM_E_STATE, "internal: a chunk's pre-pass was made for another latestBucket");
    if (apm_stats_chunk_eligible(c, n, nt)) return apm_stats_chunk(c, d, n, pp, res);  // stats only: the whole chunk in two launches
    const bool keep_any = c->cfg.flags & (APM_F_KEEP_STATS | APM_F_KEEP_FS | APM_F_ALERTS);
    if (keep_any && c->ticks_buffered + nt > c->cfg.out_ticks)
        return apm_fail(c, APM_E_WOULDBLOCK, "chunk holds more ticks than the output ring has room for: fetch results or raise out_ticks");
    const bool logs = with_pct(c);
    if (logs && pp.seg_max > c->known_max_value) c->known_max_value = pp.seg_max;
    apm_status st;
    uint64_t pos = 0;
    for (unsigned int i = 0; i < nt; ++i) {
        const TickMark tm = pp.h_ticks[i];
        if (logs) apply_seg_counts(c, pp, i);
        st = apm_launch_reduce(c, d, pos, tm.pos);  // everything that arrived before the trigger
        if (st) return st;
        st = apm_pt_append(c, d, pos, tm.pos);
        if (st) return st;
        c->pt_seq += tm.pos - pos;
        st = do_tick(c, tm.label);  // tick BEFORE the triggering transaction is added (:348-370)
        if (st) return st;
        pos = tm.pos;
    }
    if (logs) apply_seg_counts(c, pp, nt);
    st = apm_launch_reduce(c, d, pos, n);
    if (st) return st;
    st = apm_pt_append(c, d, pos, n);
    if (st) return st;
    c->pt_seq += n - pos;
    if (res) res->ticks += nt;
    return APM_OK;
}

// A call without transactions that still carries host-located ticks (a key shard that is silent while the global
// stream ticks, stream_calc_stats.js:348: the tick is triggered by ANY transaction): run them.
apm_status run_marks_only(apm_ctx* c, apm_tick_result* res) {
    if (!c->ext_armed) return APM_OK;
    for (; c->ext_next < c->ext_marks.size(); ++c->ext_next) {
        const TickMark& m = c->ext_marks[c->ext_next];
        if (m.pos != 0) return apm_fail(c, APM_E_INVALID, "tick mark beyond the call's transactions");
        apm_status st = do_tick(c, m.label);
        if (st) return st;
        if (res) res->ticks += 1;
    }
    return APM_OK;
}

// one chunk, pre-pass and ticks on the ingest stream (callers that have a single chunk)
apm_status process_chunk_device(apm_ctx* c, const apm_tx_columns& d, uint64_t n, apm_tick_result* res) {
    // The baseline of the last tick is normally launched behind the next segment's ingest kernel.  The host is
    // about to block on the tick marks and the segment table: launch it now, or the baseline stream idles.
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    apm_ctx::PrePass& pp = c->pre[0];
    c->ext_off = 0;
    c->ext_last = true;
    apm_status st = apm_prepass_begin(c, pp, d, n, c->latest, c->s_main);
    if (st) return st;
    st = apm_prepass_finish(c, pp, d, n, c->s_main);
    if (st) return st;
    return run_chunk(c, d, n, pp, res);
}

// Copies the valid records of the buffered ticks to the caller, tick-major / key-ascending.
template <typename R>
static apm_status fetch_dense(apm_ctx* c, const R* d_src, size_t per_tick, R* out, uint64_t cap, uint64_t* n_out) {
    *n_out = 0;
    const size_t total = (size_t)c->ticks_buffered * per_tick;
    if (total == 0) return APM_OK;
    if (cap < total) return apm_fail(c, APM_E_CAPACITY, "output buffer smaller than ticks_buffered * records per tick");
    // the records of the last tick are written on the baseline stream: finish that first, then copy
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaMemcpyAsync(out, d_src, total * sizeof(R), cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    c->counters.d2h_bytes += total * sizeof(R);
    size_t w = 0;
    for (size_t i = 0; i < total; ++i) {
        if (out[i].flags & 1) {
            if (w != i) out[w] = out[i];
            ++w;
        }
    }
    *n_out = w;
    return APM_OK;
}

// fills the transaction counts of a finished call from the device counters (one sync) and
// surfaces a value-log overflow as APM_E_CAPACITY
static apm_status finish_result(apm_ctx* c, const apm_counters& before, apm_tick_result* res) {
    apm_status st = refresh_counters(c);
    if (!st && c->ext_armed && c->ext_next < c->ext_marks.size())
        st = apm_fail(c, APM_E_INVALID, "tick mark beyond the call's transactions");
    if (res) {
        res->tx_ingested = c->counters.tx_ingested - before.tx_ingested;
        res->tx_dropped = c->counters.tx_dropped - before.tx_dropped;
        res->tx_invalid = c->counters.tx_invalid - before.tx_invalid;
        res->latest_bucket = c->latest;
    }
    return st;
}

// A host buffer goes through in chunks: the first ones are small so that compute starts while most of the input
// is still crossing PCIe; after that every chunk's copy hides behind the previous chunk's kernels.
// enqueue_copy(offset, count, b) puts transactions [offset, offset + count) into c->dev[b] on the copy stream.
template <typename Copy>
apm_status process_host_chunks(apm_ctx* c, uint64_t n, apm_tick_result* res, Copy enqueue_copy) {
    if (res) memset(res, 0, sizeof *res);
    const apm_counters before = c->counters;
    const uint64_t B = c->cfg.max_batch;
    std::vector<std::pair<uint64_t, uint64_t>> chunks;  // (offset, length)
    {
        // when the call spans several batches the host path is copy-bound: quarter-size chunks keep the
        // un-overlapped head (first copy) and tail (last chunk's kernels) short
        static int top_div = 0;
        if (!top_div) { const char* e = getenv("APM_CHUNK_TOP_DIV"); top_div = e ? atoi(e) : 4; if (top_div < 1) top_div = 1; }
        const uint64_t top = n > B ? std::max<uint64_t>(B / top_div, 1) : B;
        static int first_div = 0;
        if (!first_div) { const char* e = getenv("APM_CHUNK_FIRST_DIV"); first_div = e ? atoi(e) : 16; if (first_div < 1) first_div = 1; }
        uint64_t off = 0, len = n > B ? std::max<uint64_t>(B / first_div, 1) : B;
        while (off < n) {
            const uint64_t m = std::min<uint64_t>(n - off, len);
            chunks.emplace_back(off, m);
            off += m;
            static int growth = 0;  // APM_CHUNK_GROWTH_PCT: size of a chunk relative to the one before it (the link is
                                    // barely faster than the device: doubling chunks starve it during the ramp)
            if (!growth) { const char* e = getenv("APM_CHUNK_GROWTH_PCT"); growth = e ? atoi(e) : 150; if (growth < 100) growth = 100; }
            len = std::min<uint64_t>(len * growth / 100, top);
        }
    }
    const uint64_t nchunks = chunks.size();
    auto copy = [&](uint64_t ci) -> apm_status {
        const int b = (int)(ci & 1);
        CUDA_TRY(c, cudaStreamWaitEvent(c->s_copy, c->ev_consumed[b], 0));
        apm_status st = enqueue_copy(chunks[ci].first, chunks[ci].second, b);
        if (st) return st;
        CUDA_TRY(c, cudaEventRecord(c->ev_copied[b], c->s_copy));
        return APM_OK;
    };
    if (nchunks == 0) {
        apm_status st0 = run_marks_only(c, res);  // a shard that is silent for the whole call still ticks
        if (st0) return st0;
        return finish_result(c, before, res);
    }
    // Pipeline: while the ticks of chunk i are queued and run (ingest stream), chunk i + 1 is copied (copy stream)
    // and its pre-pass - tick finder, segment table, their host round trips - runs on the pre-pass stream.
    CUDA_TRY(c, cudaStreamSynchronize(c->s_pre));
    apm_status st = copy(0);
    if (st) return st;
    CUDA_TRY(c, cudaStreamWaitEvent(c->s_pre, c->ev_copied[0], 0));
    c->ext_off = 0;
    c->ext_last = nchunks == 1;
    st = apm_prepass_begin(c, c->pre[0], c->dev[0], chunks[0].second, c->latest, c->s_pre);
    if (st) return st;
    if (nchunks > 1) { st = copy(1); if (st) return st; }
    st = apm_prepass_finish(c, c->pre[0], c->dev[0], chunks[0].second, c->s_pre);
    if (st) return st;
    for (uint64_t ci = 0; ci < nchunks; ++ci) {
        const int b = (int)(ci & 1);
        const bool more = ci + 1 < nchunks;
        if (more) {  // the next chunk's tick finder starts as soon as its copy has landed
            CUDA_TRY(c, cudaStreamWaitEvent(c->s_pre, c->ev_copied[b ^ 1], 0));
            c->ext_off = chunks[ci + 1].first;
            c->ext_last = ci + 2 == nchunks;
            st = apm_prepass_begin(c, c->pre[b ^ 1], c->dev[b ^ 1], chunks[ci + 1].second, latest_after(c->pre[b]), c->s_pre);
            if (st) return st;
        }
        CUDA_TRY(c, cudaStreamWaitEvent(c->s_main, c->ev_copied[b], 0));
        st = run_chunk(c, c->dev[b], chunks[ci].second, c->pre[b], res);
        if (st == APM_E_WOULDBLOCK) { finish_result(c, before, res); return st; }  // res: what the earlier chunks consumed
        if (st) return st;
        CUDA_TRY(c, cudaEventRecord(c->ev_consumed[b], c->s_main));
        if (ci + 2 < nchunks) { st = copy(ci + 2); if (st) return st; }  // into the buffer this chunk has just released
        if (more) {
            st = apm_prepass_finish(c, c->pre[b ^ 1], c->dev[b ^ 1], chunks[ci + 1].second, c->s_pre);
            if (st) return st;
        }
    }
    return finish_result(c, before, res);
}

}  // namespace

// Launches the deferred baseline of the last tick on the baseline stream, ordered after that tick's window
// kernel and after everything queued on the ingest stream so far.
apm_status apm_flush_pending(apm_ctx* c) {
    if (!c->pending.valid) return APM_OK;
    APM_ENTER(c);
    const PendingBaseline pb = c->pending;
    c->pending.valid = false;
    CUDA_TRY(c, cudaEventRecord(c->ev_mid, c->s_main));
    CUDA_TRY(c, cudaStreamWaitEvent(c->s_base, c->ev_mid, 0));
    apm_status st = launch_baseline_group(c, pb, c->s_base);
    if (st) return st;
    CUDA_TRY(c, cudaEventRecord(c->ev_base[pb.ev], c->s_base));
    return APM_OK;
}

cudaError_t apm_preload_capi(); cudaError_t apm_preload_comm(); cudaError_t apm_preload_k1_ingest(); cudaError_t apm_preload_k2_tick();
cudaError_t apm_preload_k3_onesweep(); cudaError_t apm_preload_k3_percentile_op(); cudaError_t apm_preload_k3_sort();
cudaError_t apm_preload_k4_baseline(); cudaError_t apm_preload_k5_passthrough(); cudaError_t apm_preload_k6_logparse();
cudaError_t apm_preload_k3_groupsort(); cudaError_t apm_preload_k1_chunk();

extern "C" {

// CUDA loads a kernel when it is first launched, and that load can need the device to drain.  A kernel that spins on
// another GPU's progress (an NCCL collective waiting for its peer) never drains, so two ranks that each meet a
// not-yet-loaded kernel while the other's collective is in flight wait for each other for ever.  Loading every kernel
// up front removes the first-use loads from the steady state (apm_comm_init does it for its device).
apm_status apm_preload_kernels(int32_t device) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return apm_fail(nullptr, APM_E_DEVICE, "apm_preload_kernels: no such CUDA device");
    ApmDeviceGuard guard(device);
    static unsigned long long done = 0;
    if (!apm_first_on_device(&done)) return APM_OK;
    cudaError_t (*const parts[])() = {apm_preload_capi, apm_preload_comm, apm_preload_k1_ingest, apm_preload_k1_chunk, apm_preload_k2_tick, apm_preload_k3_onesweep,
                                      apm_preload_k3_percentile_op, apm_preload_k3_sort, apm_preload_k3_groupsort, apm_preload_k4_baseline,
                                      apm_preload_k5_passthrough, apm_preload_k6_logparse};
    for (auto part : parts) {
        const cudaError_t e = part();
        if (e != cudaSuccess) { done = 0; return apm_fail_cuda(nullptr, e, "cudaFuncGetAttributes (kernel preload)", __FILE__, __LINE__); }
    }
    return APM_OK;
}

uint32_t apm_abi_version(void) { return APM_ABI_VERSION; }

const char* apm_last_error(const apm_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

double apm_quantize(double x, int32_t digits) { return apm_quantize_fixed(x, digits == 2 ? 2 : 1); }

void apm_ctx_destroy(apm_ctx* c) {
    if (!c) return;
    APM_ENTER(c);
    if (c->s_main) cudaStreamSynchronize(c->s_main);
    c->pending.valid = false;
    if (c->s_base) cudaStreamSynchronize(c->s_base);
    if (c->s_copy) cudaStreamSynchronize(c->s_copy);
    if (c->s_pre) cudaStreamSynchronize(c->s_pre);
    cudaFree(c->acc_raw); cudaFree(c->seen); cudaFree(c->slot_meta);
    for (int i = 0; i <= c->S && i <= APM_MAX_SLOTS; ++i) cudaFree(c->log_buf[i]);
    cudaFree(c->d_log_ptrs); cudaFree(c->sort_hist); cudaFree(c->sort_base); cudaFree(c->sort_scratch); cudaFree(c->gs_scratch);
    for (int i = 0; i < 2; ++i) {
        apm_ctx::PrePass& pp = c->pre[i];
        cudaFree(pp.tf_tile_max); cudaFree(pp.tf_tile_prefix); cudaFree(pp.d_ticks); cudaFree(pp.d_nticks);
        cudaFreeHost(pp.h_ticks); cudaFreeHost(pp.h_nticks);
        cudaFree(pp.d_seg_counts); cudaFreeHost(pp.h_seg_counts);
        if (pp.ev) cudaEventDestroy(pp.ev);
    }
    apm_lp_destroy(c);
    cudaFree(c->pk[0]); cudaFree(c->pk[1]); cudaFree(c->esc[0]); cudaFree(c->esc[1]);
    cudaFreeHost(c->pin.service_id); cudaFreeHost(c->pin.jvm_id); cudaFreeHost(c->pin.end_ts_ms); cudaFreeHost(c->pin.elapsed_ms);
    free_columns_device(&c->dev[0]); free_columns_device(&c->dev[1]);
    cudaFree(c->d_stats); cudaFree(c->d_fs); cudaFree(c->d_alerts); cudaFree(c->d_alert_count); cudaFree(c->d_alert_blocks); cudaFree(c->zin[0]); cudaFree(c->zin[1]); cudaFree(c->emit[0]); cudaFree(c->emit[1]);
    for (int l = 0; l < APM_MAX_LAGS; ++l) { cudaFree(c->ring[l]); cudaFree(c->ring_len[l]); cudaFree(c->ring_next[l]);
        cudaFree(c->inc_sum[l]); cudaFree(c->inc_comp[l]); cudaFree(c->inc_cnt[l]); cudaFree(c->inc_nz[l]); cudaFree(c->thr[l]); cudaFree(c->infl[l]); }
    cudaFree(c->d_counters); cudaFreeHost(c->h_counters); cudaFree(c->chunk_planes); cudaFree(c->chunk_first_seg); cudaFree(c->chunk_desc);
    cudaFree(c->alert_count); cudaFree(c->alert_hard_max); cudaFree(c->alert_suppressed);
    cudaFree(c->pt_pend[0]); cudaFree(c->pt_pend[1]); cudaFree(c->pt_emit); cudaFree(c->pt_tmp); cudaFree(c->pt_out); cudaFree(c->pt_cursors);
    for (int i = 0; i < 2; ++i) { if (c->ev_copied[i]) cudaEventDestroy(c->ev_copied[i]); if (c->ev_consumed[i]) cudaEventDestroy(c->ev_consumed[i]); }
    for (int i = 0; i < 4; ++i) { if (c->ev_stats[i]) cudaEventDestroy(c->ev_stats[i]); if (c->ev_base[i]) cudaEventDestroy(c->ev_base[i]); }
    if (c->ev_mid) cudaEventDestroy(c->ev_mid);
    if (c->ev_d2h) cudaEventDestroy(c->ev_d2h);
    if (c->s_d2h) { cudaStreamSynchronize(c->s_d2h); cudaStreamDestroy(c->s_d2h); }
    if (c->s_base) cudaStreamDestroy(c->s_base);
    if (c->s_main) cudaStreamDestroy(c->s_main);
    if (c->s_copy) cudaStreamDestroy(c->s_copy);
    if (c->s_pre) cudaStreamDestroy(c->s_pre);
    apm_comm_destroy_internal(c);
    for (auto& sp : c->prof_spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (auto e : c->prof_pool) cudaEventDestroy(e);
    for (int i = 0; i < 2; ++i) if (c->timer[i]) cudaEventDestroy(c->timer[i]);
    delete c;
}

apm_status apm_ctx_create(const apm_config* cfg, apm_ctx** out) {
    if (!out) return apm_fail(nullptr, APM_E_INVALID, "out is NULL");
    *out = nullptr;
    apm_status st = validate(cfg);
    if (st) return st;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return apm_fail(nullptr, APM_E_DEVICE, std::string("no CUDA device available (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (cfg->device < 0 || cfg->device >= ndev) return apm_fail(nullptr, APM_E_INVALID, "device ordinal out of range");
    apm_ctx* c = new apm_ctx();
    c->cfg = *cfg;
    memset(&c->counters, 0, sizeof c->counters);
    for (int i = 0; i <= APM_MAX_SLOTS; ++i) c->log_buf[i] = nullptr;
    for (int i = 0; i < APM_MAX_SLOTS; ++i) c->slot_label[i] = -1;
    c->K = cfg->num_services * cfg->num_jvms;
    c->keep = cfg->window_buckets + cfg->buffer_buckets;
    c->S = c->keep + 1;
    c->inv_bucket = 1.0 / (double)cfg->bucket_ms;
    if (const char* e = getenv("APM_K4_STAGES")) c->k4_stages = atoi(e);
    if (const char* e = getenv("APM_K4_LDG")) c->k4_force_ldg = atoi(e) != 0;
    if (const char* e = getenv("APM_K4_FEW")) c->k4_few = atoi(e) != 0;
    if (const char* e = getenv("APM_K4_FEW_WAVES")) c->k4_few_waves = std::max(1, atoi(e));
    if (getenv("APM_SMEM_CARVEOUT")) apm_preload_kernels(c->cfg.device);  // the carve-out preference is applied by the preload
    if (const char* e = getenv("APM_K1_TPC")) { int v = atoi(e); if (v > 0) c->k1_tiles_per_cta = (uint32_t)v; }
    // windowSz * intervalLengthSec / 60.0 (stream_calc_stats.js:186)
    c->tpm_div = (double)cfg->window_buckets * ((double)cfg->bucket_ms / 1000.0) / 60.0;
    for (int i = 0; i < cfg->num_percentiles; ++i) c->pfrac[i] = cfg->percentiles[i] / 100.0;  // util_methods.js:129
    auto fail = [&](apm_status s) { g_create_error = c->err; apm_ctx_destroy(c); return s; };
#define TRY_CREATE(expr) do { cudaError_t e2_ = (expr); if (e2_ != cudaSuccess) { apm_fail_cuda(c, e2_, #expr, __FILE__, __LINE__); return fail(APM_E_DEVICE); } } while (0)
    APM_ENTER(c);  // the caller's current device is restored when create returns
    TRY_CREATE(cudaSetDevice(cfg->device));
    c->num_sms = apm_sm_count();
    {
        // the ingest / sort chain is the latency-critical one: its CTAs go first when both streams have work
        int prio_lo = 0, prio_hi = 0;
        TRY_CREATE(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        TRY_CREATE(cudaStreamCreateWithPriority(&c->s_main, cudaStreamNonBlocking, prio_hi));
        TRY_CREATE(cudaStreamCreateWithPriority(&c->s_copy, cudaStreamNonBlocking, prio_hi));
        TRY_CREATE(cudaStreamCreateWithPriority(&c->s_base, cudaStreamNonBlocking, prio_lo));
        TRY_CREATE(cudaStreamCreateWithPriority(&c->s_pre, cudaStreamNonBlocking, prio_hi));
    }
    for (int i = 0; i < 4; ++i) {
        TRY_CREATE(cudaEventCreateWithFlags(&c->ev_stats[i], cudaEventDisableTiming));
        TRY_CREATE(cudaEventCreateWithFlags(&c->ev_base[i], cudaEventDisableTiming));
    }
    TRY_CREATE(cudaEventCreateWithFlags(&c->ev_mid, cudaEventDisableTiming));
    TRY_CREATE(cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
    TRY_CREATE(cudaEventCreateWithFlags(&c->ev_d2h, cudaEventDisableTiming));
    if (const char* e = getenv("APM_NO_OVERLAP")) c->overlap = atoi(e) == 0;
    if (const char* e = getenv("APM_STATS_CHUNK")) c->stats_chunk = atoi(e) != 0;
    if (const char* e = getenv("APM_STATS_CHUNK_MIN_N")) c->stats_chunk_min_n = strtoull(e, nullptr, 10);
    if (const char* e = getenv("APM_STATS_CHUNK_MIN_SEG")) c->stats_chunk_min_seg = strtoull(e, nullptr, 10);
    if (const char* e = getenv("APM_STATS_CHUNK_DIRECT")) c->stats_chunk_direct = atoi(e) != 0;
    for (int i = 0; i < 2; ++i) {
        TRY_CREATE(cudaEventCreateWithFlags(&c->ev_copied[i], cudaEventDisableTiming));
        TRY_CREATE(cudaEventCreateWithFlags(&c->ev_consumed[i], cudaEventDisableTiming));
    }
    const size_t K = c->K;
    {
        const size_t n = (size_t)c->S * K;
        TRY_CREATE(cudaMalloc(&c->acc_raw, n * APM_ACC_BYTES_PER_ENTRY));
        char* b = (char*)c->acc_raw;
        c->acc.sum = (unsigned long long*)b;
        c->acc.sq_lo = c->acc.sum + n;
        c->acc.sq_hi = c->acc.sq_lo + n;
        c->acc.cnt = (uint32_t*)(c->acc.sq_hi + n);
        c->acc.vmin = c->acc.cnt + n;
        c->acc.vmax = c->acc.vmin + n;
        c->acc.run_start = c->acc.vmax + n;
    }
    TRY_CREATE(dalloc(&c->seen, K));
    TRY_CREATE(cudaMemsetAsync(c->seen, 0, K, c->s_main));
    TRY_CREATE(dalloc(&c->slot_meta, (size_t)c->S));
    TRY_CREATE(cudaMemsetAsync(c->slot_meta, 0, sizeof(SlotMeta) * c->S, c->s_main));
    TRY_CREATE(dalloc(&c->d_log_ptrs, (size_t)APM_MAX_SLOTS));
    if (cfg->flags & APM_F_PASSTHROUGH) {
        if (cfg->max_tx_per_bucket == 0) { c->err = "max_tx_per_bucket must be > 0 with APM_F_PASSTHROUGH"; return fail(APM_E_INVALID); }
        c->pt_cap = cfg->max_tx_per_bucket * (uint64_t)(cfg->buffer_buckets + 3);  // due at buffer+1 buckets of age, plus late arrivals
        c->pt_out_cap = 2 * c->pt_cap;
        TRY_CREATE(dalloc(&c->pt_pend[0], (size_t)c->pt_cap));
        TRY_CREATE(dalloc(&c->pt_pend[1], (size_t)c->pt_cap));
        TRY_CREATE(dalloc(&c->pt_emit, (size_t)c->pt_cap));
        TRY_CREATE(dalloc(&c->pt_tmp, (size_t)c->pt_cap));
        TRY_CREATE(dalloc(&c->pt_out, (size_t)c->pt_out_cap));
        TRY_CREATE(dalloc(&c->pt_cursors, (size_t)3));
        TRY_CREATE(cudaMemsetAsync(c->pt_cursors, 0, 3 * sizeof(unsigned long long), c->s_main));
        c->pt_tick_counts.assign(cfg->out_ticks, 0);
    }
    if (with_pct(c) || c->pt_cap) {
        const uint64_t sort_cap = std::max<uint64_t>(with_pct(c) ? cfg->max_tx_per_bucket : 0, c->pt_cap);
        c->sort_tiles_cap = (sort_cap + 4095) / 4096;
        TRY_CREATE(dalloc(&c->sort_hist, (size_t)256 * c->sort_tiles_cap));
        TRY_CREATE(dalloc(&c->sort_base, (size_t)256));
    }
    if (with_pct(c)) {
        c->log_cap = cfg->max_tx_per_bucket;
        for (int i = 0; i <= c->S; ++i) TRY_CREATE(dalloc(&c->log_buf[i], (size_t)c->log_cap));
        c->sort_scratch_bytes = apm_onesweep_scratch_bytes(c->log_cap);
        if (c->sort_scratch_bytes <= ((size_t)1 << 30)) TRY_CREATE(cudaMalloc((void**)&c->sort_scratch, c->sort_scratch_bytes));
        else c->sort_scratch_bytes = 0;  // very large logs keep the three-kernel sort
        if (const char* e = getenv("APM_SORT_LEGACY")) c->sort_legacy = atoi(e) != 0;
        if (const char* e = getenv("APM_SORT_GROUP")) c->sort_group = atoi(e) != 0;
        if (const char* e = getenv("APM_SMALL_TICK")) c->small_tick = atoi(e) != 0;
        TRY_CREATE(cudaMalloc((void**)&c->gs_scratch, apm_groupsort_scratch_bytes(c->K)));
    }
    TRY_CREATE(cudaMemcpyAsync(c->d_log_ptrs, c->log_buf, sizeof(uint64_t*) * APM_MAX_SLOTS, cudaMemcpyHostToDevice, c->s_main));
    c->tf_tiles_cap = (cfg->max_batch + 4095) / 4096;
    for (int i = 0; i < 2; ++i) {
        apm_ctx::PrePass& pp = c->pre[i];
        const size_t cells = (size_t)(c->max_ticks_per_chunk + 1) * (c->keep + 1) + 1;
        TRY_CREATE(dalloc(&pp.tf_tile_max, (size_t)c->tf_tiles_cap));
        TRY_CREATE(dalloc(&pp.tf_tile_prefix, (size_t)c->tf_tiles_cap));
        TRY_CREATE(dalloc(&pp.d_ticks, (size_t)c->max_ticks_per_chunk));
        TRY_CREATE(dalloc(&pp.d_nticks, (size_t)1));
        TRY_CREATE(cudaMallocHost((void**)&pp.h_ticks, sizeof(TickMark) * c->max_ticks_per_chunk));
        TRY_CREATE(cudaMallocHost((void**)&pp.h_nticks, sizeof(unsigned int)));
        TRY_CREATE(dalloc(&pp.d_seg_counts, cells));
        TRY_CREATE(cudaMallocHost((void**)&pp.h_seg_counts, sizeof(uint32_t) * cells));
        TRY_CREATE(cudaEventCreateWithFlags(&pp.ev, cudaEventDisableTiming));
    }
    for (int s = 0; s < APM_MAX_SLOTS; ++s) c->host_count[s] = c->host_sorted[s] = 0;
    for (int i = 0; i < 2; ++i) {
        TRY_CREATE(dalloc(&c->zin[i], 4 * K));
        TRY_CREATE(dalloc(&c->emit[i], K));
        TRY_CREATE(cudaMemsetAsync(c->emit[i], 0, K, c->s_main));
    }
    if (cfg->flags & APM_F_KEEP_STATS) TRY_CREATE(dalloc(&c->d_stats, (size_t)cfg->out_ticks * K));
    if (with_z(c)) {
        const size_t fs_ticks = (cfg->flags & APM_F_KEEP_FS) ? cfg->out_ticks : 1;
        TRY_CREATE(dalloc(&c->d_fs, fs_ticks * cfg->num_lags * K));
        if (cfg->flags & APM_F_ALERTS) {
            if (cfg->max_alerts_per_tick == 0) { c->err = "max_alerts_per_tick must be > 0 with APM_F_ALERTS"; return fail(APM_E_INVALID); }
            TRY_CREATE(dalloc(&c->d_alerts, (size_t)cfg->out_ticks * cfg->max_alerts_per_tick));
            TRY_CREATE(dalloc(&c->d_alert_count, (size_t)cfg->out_ticks));
            TRY_CREATE(dalloc(&c->d_alert_blocks, (size_t)K * cfg->num_lags / 256 + 3));
            TRY_CREATE(cudaMemsetAsync(c->d_alert_blocks, 0, sizeof(uint32_t) * ((size_t)K * cfg->num_lags / 256 + 3), c->s_main));
            c->k4_done = c->d_alert_blocks + ((size_t)K * cfg->num_lags / 256 + 2);  // CTAs-finished counter of the few-list walk
            TRY_CREATE(cudaMemsetAsync(c->d_alert_count, 0, sizeof(unsigned int) * cfg->out_ticks, c->s_main));
        }
        c->ring_pitch = (3 * K + 31) & ~(size_t)31;
        TRY_CREATE(dalloc(&c->alert_count, (size_t)cfg->num_lags * K));
        TRY_CREATE(cudaMemsetAsync(c->alert_count, 0, sizeof(uint32_t) * cfg->num_lags * K, c->s_main));
        TRY_CREATE(dalloc(&c->alert_hard_max, K));
        TRY_CREATE(cudaMemsetAsync(c->alert_hard_max, 0, sizeof(double) * K, c->s_main));
        TRY_CREATE(dalloc(&c->alert_suppressed, K));
        TRY_CREATE(cudaMemsetAsync(c->alert_suppressed, 0, K, c->s_main));
        for (int l = 0; l < cfg->num_lags; ++l) {
            const size_t L = (size_t)cfg->lags[l];
            TRY_CREATE(dalloc(&c->ring[l], L * c->ring_pitch));
            TRY_CREATE(cudaMemsetAsync(c->ring[l], 0, L * c->ring_pitch * sizeof(double), c->s_main));
            TRY_CREATE(dalloc(&c->ring_len[l], 3 * K));
            TRY_CREATE(cudaMemsetAsync(c->ring_len[l], 0, sizeof(uint32_t) * 3 * K, c->s_main));
            if (cfg->flags & APM_F_INCREMENTAL) {
                TRY_CREATE(dalloc(&c->inc_sum[l], 3 * K)); TRY_CREATE(cudaMemsetAsync(c->inc_sum[l], 0, sizeof(double) * 3 * K, c->s_main));
                TRY_CREATE(dalloc(&c->inc_comp[l], 3 * K)); TRY_CREATE(cudaMemsetAsync(c->inc_comp[l], 0, sizeof(double) * 3 * K, c->s_main));
                TRY_CREATE(dalloc(&c->inc_cnt[l], 3 * K)); TRY_CREATE(cudaMemsetAsync(c->inc_cnt[l], 0, sizeof(uint32_t) * 3 * K, c->s_main));
                TRY_CREATE(dalloc(&c->inc_nz[l], 3 * K)); TRY_CREATE(cudaMemsetAsync(c->inc_nz[l], 0, sizeof(uint32_t) * 3 * K, c->s_main));
            }
            TRY_CREATE(dalloc(&c->ring_next[l], 3 * K));
            TRY_CREATE(cudaMemsetAsync(c->ring_next[l], 0, sizeof(uint32_t) * 3 * K, c->s_main));
            TRY_CREATE(dalloc(&c->thr[l], K));
            TRY_CREATE(dalloc(&c->infl[l], K));
            std::vector<double> t(K, cfg->default_threshold[l]), f(K, cfg->default_influence[l]);
            TRY_CREATE(cudaMemcpy(c->thr[l], t.data(), K * sizeof(double), cudaMemcpyHostToDevice));
            TRY_CREATE(cudaMemcpy(c->infl[l], f.data(), K * sizeof(double), cudaMemcpyHostToDevice));
        }
    }
    if (with_z(c)) { st = apm_baseline_make_tmaps(c); if (st) return fail(st); }
    TRY_CREATE(dalloc(&c->d_counters, (size_t)1));
    TRY_CREATE(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->s_main));
    TRY_CREATE(cudaMallocHost((void**)&c->h_counters, sizeof(DevCounters)));
    for (int i = 0; i < 2; ++i) {
        st = alloc_columns_device(c, &c->dev[i], cfg->max_batch);
        if (st) return fail(st);
    }
    TRY_CREATE(cudaMallocHost((void**)&c->pin.service_id, cfg->max_batch * sizeof(uint32_t)));
    TRY_CREATE(cudaMallocHost((void**)&c->pin.jvm_id, cfg->max_batch * sizeof(uint32_t)));
    TRY_CREATE(cudaMallocHost((void**)&c->pin.end_ts_ms, cfg->max_batch * sizeof(int64_t)));
    TRY_CREATE(cudaMallocHost((void**)&c->pin.elapsed_ms, cfg->max_batch * sizeof(int32_t)));
    TRY_CREATE(cudaStreamSynchronize(c->s_main));
#undef TRY_CREATE
    *out = c;
    return APM_OK;
}

apm_status apm_sync(apm_ctx* c) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    return sync_all(c);
}

apm_status apm_ingest_acquire(apm_ctx* c, uint64_t n, apm_tx_columns* out) {
    if (!c || !out) return APM_E_INVALID;
    APM_ENTER(c);
    if (n > c->cfg.max_batch) return apm_fail(c, APM_E_CAPACITY, "acquire larger than max_batch");
    // the previous commit's copy must have left the pinned slices
    CUDA_TRY(c, cudaStreamSynchronize(c->s_copy));
    *out = c->pin;
    c->acquired = n;
    return APM_OK;
}

apm_status apm_ingest_commit(apm_ctx* c, uint64_t n) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (n > c->acquired) return apm_fail(c, APM_E_STATE, "commit without a matching acquire");
    c->acquired = 0;
    if (n == 0) return APM_OK;
    if (c->latest == 0) return apm_fail(c, APM_E_STATE, "commit before the first tick: run apm_tick for the first bucket label");
    apm_tx_columns& d = c->dev[0];
    CUDA_TRY(c, cudaStreamWaitEvent(c->s_copy, c->ev_consumed[0], 0));
    CUDA_TRY(c, cudaMemcpyAsync(d.service_id, c->pin.service_id, n * 4, cudaMemcpyHostToDevice, c->s_copy));
    CUDA_TRY(c, cudaMemcpyAsync(d.jvm_id, c->pin.jvm_id, n * 4, cudaMemcpyHostToDevice, c->s_copy));
    CUDA_TRY(c, cudaMemcpyAsync(d.end_ts_ms, c->pin.end_ts_ms, n * 8, cudaMemcpyHostToDevice, c->s_copy));
    CUDA_TRY(c, cudaMemcpyAsync(d.elapsed_ms, c->pin.elapsed_ms, n * 4, cudaMemcpyHostToDevice, c->s_copy));
    CUDA_TRY(c, cudaEventRecord(c->ev_copied[0], c->s_copy));
    CUDA_TRY(c, cudaStreamWaitEvent(c->s_main, c->ev_copied[0], 0));
    c->counters.h2d_bytes += n * 20;
    apm_status st;
    if (with_pct(c)) {  // one segment under the current latestBucket
        st = apm_prepass_segment(c, c->pre[0], d, n, c->latest, c->s_main);
        if (st) return st;
        if (c->pre[0].seg_max > c->known_max_value) c->known_max_value = c->pre[0].seg_max;
        apply_seg_counts(c, c->pre[0], 0);
    }
    st = apm_launch_reduce(c, d, 0, n);
    if (st) return st;
    st = apm_pt_append(c, d, 0, n);
    if (st) return st;
    c->pt_seq += n;
    CUDA_TRY(c, cudaEventRecord(c->ev_consumed[0], c->s_main));
    return APM_OK;
}

apm_status apm_tick(apm_ctx* c, int64_t latest_bucket, apm_tick_result* res) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    apm_status st = do_tick(c, latest_bucket);
    if (st) return st;
    if (res) {
        memset(res, 0, sizeof *res);
        res->ticks = 1;
        res->latest_bucket = c->latest;
    }
    return APM_OK;
}

apm_status apm_process_device(apm_ctx* c, const apm_tx_columns* cols, uint64_t n, apm_tick_result* res) {
    if (!c || !cols) return APM_E_INVALID;
    APM_ENTER(c);
    ExtMarksScope ext_scope_{c};
    if (res) memset(res, 0, sizeof *res);
    const apm_counters before = c->counters;
    const uint64_t B = c->cfg.max_batch;
    if (n <= B) {
        apm_status st = n ? process_chunk_device(c, *cols, n, res) : run_marks_only(c, res);
        if (st) return st;
        return finish_result(c, before, res);
    }
    // several chunks: the pre-pass of chunk i + 1 runs on its own stream while the ticks of chunk i are queued
    auto chunk = [&](uint64_t ci, uint64_t* m) {
        const uint64_t off = ci * B;
        *m = std::min<uint64_t>(n - off, B);
        return apm_tx_columns{cols->service_id + off, cols->jvm_id + off, cols->end_ts_ms + off, cols->elapsed_ms + off};
    };
    const uint64_t nchunks = (n + B - 1) / B;
    CUDA_TRY(c, cudaStreamSynchronize(c->s_pre));
    uint64_t m = 0, m1 = 0;
    apm_tx_columns d = chunk(0, &m);
    c->ext_off = 0;
    c->ext_last = false;
    apm_status st = apm_prepass_begin(c, c->pre[0], d, m, c->latest, c->s_pre);
    if (st) return st;
    st = apm_prepass_finish(c, c->pre[0], d, m, c->s_pre);
    if (st) return st;
    for (uint64_t ci = 0; ci < nchunks; ++ci) {
        const int b = (int)(ci & 1);
        const bool more = ci + 1 < nchunks;
        apm_tx_columns d1 = d;
        if (more) {
            d1 = chunk(ci + 1, &m1);
            c->ext_off = (ci + 1) * B;
            c->ext_last = ci + 2 == nchunks;
            st = apm_prepass_begin(c, c->pre[b ^ 1], d1, m1, latest_after(c->pre[b]), c->s_pre);
            if (st) return st;
        }
        st = run_chunk(c, d, m, c->pre[b], res);
        if (st == APM_E_WOULDBLOCK) { finish_result(c, before, res); return st; }  // res: what the earlier chunks consumed
        if (st) return st;
        if (more) {
            st = apm_prepass_finish(c, c->pre[b ^ 1], d1, m1, c->s_pre);
            if (st) return st;
            d = d1;
            m = m1;
        }
    }
    return finish_result(c, before, res);
}

apm_status apm_process_host(apm_ctx* c, const apm_tx_columns* h, uint64_t n, apm_tick_result* res) {
    if (!c || !h) return APM_E_INVALID;
    APM_ENTER(c);
    ExtMarksScope ext_scope_{c};
    return process_host_chunks(c, n, res, [&](uint64_t off, uint64_t m, int b) -> apm_status {
        apm_tx_columns& d = c->dev[b];
        CUDA_TRY(c, cudaMemcpyAsync(d.service_id, h->service_id + off, m * 4, cudaMemcpyHostToDevice, c->s_copy));
        CUDA_TRY(c, cudaMemcpyAsync(d.jvm_id, h->jvm_id + off, m * 4, cudaMemcpyHostToDevice, c->s_copy));
        CUDA_TRY(c, cudaMemcpyAsync(d.end_ts_ms, h->end_ts_ms + off, m * 8, cudaMemcpyHostToDevice, c->s_copy));
        CUDA_TRY(c, cudaMemcpyAsync(d.elapsed_ms, h->elapsed_ms + off, m * 4, cudaMemcpyHostToDevice, c->s_copy));
        c->counters.h2d_bytes += m * 20;
        return APM_OK;
    });
}

apm_status apm_process_host_packed(apm_ctx* c, const apm_tx_packed* h, int64_t base_ts_ms, uint64_t n, apm_tick_result* res) {
    if (!c || !h) return APM_E_INVALID;
    APM_ENTER(c);
    ExtMarksScope ext_scope_{c};
    if (base_ts_ms < 0) return apm_fail(c, APM_E_INVALID, "process_host_packed: base_ts_ms must not be negative");
    if (!c->pk[0]) {  // 12-byte staging, one third of the wide staging, created at the first packed call
        for (int i = 0; i < 2; ++i) CUDA_TRY(c, cudaMalloc((void**)&c->pk[i], c->cfg.max_batch * 3 * sizeof(uint32_t)));
    }
    return process_host_chunks(c, n, res, [&](uint64_t off, uint64_t m, int b) -> apm_status {
        uint32_t* p = c->pk[b];
        const uint64_t B = c->cfg.max_batch;
        CUDA_TRY(c, cudaMemcpyAsync(p, h->key + off, m * 4, cudaMemcpyHostToDevice, c->s_copy));
        CUDA_TRY(c, cudaMemcpyAsync(p + B, h->end_ts_delta_ms + off, m * 4, cudaMemcpyHostToDevice, c->s_copy));
        CUDA_TRY(c, cudaMemcpyAsync(p + 2 * B, h->elapsed_ms + off, m * 4, cudaMemcpyHostToDevice, c->s_copy));
        widen_tx_kernel<<<(unsigned)((m + 255) / 256), 256, 0, c->s_copy>>>(p, p + B, p + 2 * B, base_ts_ms, m, c->dev[b]);
        c->launches += 1;
        CUDA_TRY(c, cudaGetLastError());
        c->counters.h2d_bytes += m * 12;
        return APM_OK;
    });
}

apm_status apm_process_host_packed8(apm_ctx* c, const uint64_t* tx, const apm_tx_escape* escapes, uint32_t num_escapes,
                                    int64_t base_ts_ms, uint64_t n, apm_tick_result* res) {
    if (!c || (!tx && n) || (!escapes && num_escapes)) return APM_E_INVALID;
    APM_ENTER(c);
    ExtMarksScope ext_scope_{c};
    if (base_ts_ms < 0) return apm_fail(c, APM_E_INVALID, "process_host_packed8: base_ts_ms must not be negative");
    if (n > 0xFFFFFFFFull) return apm_fail(c, APM_E_CAPACITY, "process_host_packed8: at most 2^32 - 1 transactions per call (escape indices are 32-bit)");
    for (uint32_t i = 1; i < num_escapes; ++i)
        if (escapes[i].index <= escapes[i - 1].index) return apm_fail(c, APM_E_INVALID, "process_host_packed8: escapes must be ascending by index");
    if (!c->pk[0]) {  // shared with the 12-byte form: [3][max_batch] u32 holds max_batch u64 words
        for (int i = 0; i < 2; ++i) CUDA_TRY(c, cudaMalloc((void**)&c->pk[i], c->cfg.max_batch * 3 * sizeof(uint32_t)));
    }
    return process_host_chunks(c, n, res, [&](uint64_t off, uint64_t m, int b) -> apm_status {
        uint64_t* p = reinterpret_cast<uint64_t*>(c->pk[b]);
        CUDA_TRY(c, cudaMemcpyAsync(p, tx + off, m * 8, cudaMemcpyHostToDevice, c->s_copy));
        // the chunk's slice of the escape list
        const apm_tx_escape* lo = std::lower_bound(escapes, escapes + num_escapes, off, [](const apm_tx_escape& e, uint64_t v) { return e.index < v; });
        const apm_tx_escape* hi = std::lower_bound(lo, escapes + num_escapes, off + m, [](const apm_tx_escape& e, uint64_t v) { return e.index < v; });
        const uint32_t ne = (uint32_t)(hi - lo);
        if (ne > c->esc_cap[b]) {
            CUDA_TRY(c, cudaStreamSynchronize(c->s_copy));
            cudaFree(c->esc[b]);
            c->esc[b] = nullptr;
            c->esc_cap[b] = std::max<uint32_t>(2 * ne, 1024);
            CUDA_TRY(c, cudaMalloc((void**)&c->esc[b], (size_t)c->esc_cap[b] * sizeof(apm_tx_escape)));
        }
        if (ne) CUDA_TRY(c, cudaMemcpyAsync(c->esc[b], lo, (size_t)ne * sizeof(apm_tx_escape), cudaMemcpyHostToDevice, c->s_copy));
        widen_tx8_kernel<<<(unsigned)((m + 255) / 256), 256, 0, c->s_copy>>>(p, c->esc[b], ne, off, base_ts_ms, m, c->cfg.num_services, c->dev[b]);
        c->launches += 1;
        CUDA_TRY(c, cudaGetLastError());
        c->counters.h2d_bytes += m * 8 + (uint64_t)ne * sizeof(apm_tx_escape);
        return APM_OK;
    });
}

apm_status apm_set_tick_marks(apm_ctx* c, const apm_tick_mark* marks, uint32_t n) {
    if (!c || (!marks && n)) return APM_E_INVALID;
    c->ext_marks.clear();
    c->ext_next = 0;
    int64_t prev = c->latest;
    uint64_t pos = 0;
    for (uint32_t i = 0; i < n; ++i) {
        if (marks[i].pos < pos || marks[i].label <= prev) {
            c->ext_armed = false;
            c->ext_marks.clear();
            return apm_fail(c, APM_E_INVALID, "tick marks must be ascending in position and strictly ascending in label, above latestBucket");
        }
        pos = marks[i].pos;
        prev = marks[i].label;
        c->ext_marks.push_back(TickMark{(unsigned long long)marks[i].pos, (long long)marks[i].label});
    }
    c->ext_armed = true;  // even with n == 0: the next call then runs without any tick
    return APM_OK;
}

apm_status apm_set_key_params(apm_ctx* c, uint32_t key_lo, uint32_t n, int32_t l, const double* thr, const double* infl) {
    if (!c || !thr || !infl) return APM_E_INVALID;
    APM_ENTER(c);
    if (!with_z(c) || l < 0 || l >= c->cfg.num_lags || (uint64_t)key_lo + n > c->K) return apm_fail(c, APM_E_INVALID, "set_key_params: range");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    CUDA_TRY(c, cudaMemcpy(c->thr[l] + key_lo, thr, n * sizeof(double), cudaMemcpyHostToDevice));
    CUDA_TRY(c, cudaMemcpy(c->infl[l] + key_lo, infl, n * sizeof(double), cudaMemcpyHostToDevice));
    return APM_OK;
}

apm_status apm_set_alert_rules(apm_ctx* c, const apm_alert_rules* r) {
    if (!c || !r) return APM_E_INVALID;
    APM_ENTER(c);
    if (r->struct_size != sizeof(apm_alert_rules)) return apm_fail(c, APM_E_INVALID, "apm_alert_rules.struct_size mismatch");
    if (!with_z(c)) return apm_fail(c, APM_E_STATE, "alert rules need APM_F_ZSCORE and at least one lag");
    if (r->enabled && !(c->cfg.flags & (APM_F_KEEP_FS | APM_F_ALERTS))) return apm_fail(c, APM_E_STATE, "alert rules need APM_F_KEEP_FS or APM_F_ALERTS");
    apm_status st = sync_all(c);
    if (st) return st;
    c->alert_rules = *r;
    return APM_OK;
}

apm_status apm_set_alert_key_params(apm_ctx* c, uint32_t key_lo, uint32_t n, const double* hard_max_ms, const uint8_t* suppressed) {
    if (!c || !hard_max_ms || !suppressed) return APM_E_INVALID;
    APM_ENTER(c);
    if (!with_z(c) || (uint64_t)key_lo + n > c->K) return apm_fail(c, APM_E_INVALID, "set_alert_key_params: range");
    apm_status st = sync_all(c);
    if (st) return st;
    CUDA_TRY(c, cudaMemcpy(c->alert_hard_max + key_lo, hard_max_ms, n * sizeof(double), cudaMemcpyHostToDevice));
    CUDA_TRY(c, cudaMemcpy(c->alert_suppressed + key_lo, suppressed, n, cudaMemcpyHostToDevice));
    return APM_OK;
}

apm_status apm_drop_lag(apm_ctx* c, int32_t l) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (!with_z(c) || l < 0 || l >= c->cfg.num_lags) return apm_fail(c, APM_E_INVALID, "drop_lag: index");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaMemsetAsync(c->ring_len[l], 0, sizeof(uint32_t) * 3 * c->K, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    c->inc_dirty[l] = true;
    return APM_OK;
}

apm_status apm_fetch_stats(apm_ctx* c, apm_stat_record* out, uint64_t cap, uint64_t* n) {
    if (!c || !n) return APM_E_INVALID;
    APM_ENTER(c);
    if (!(c->cfg.flags & APM_F_KEEP_STATS)) return apm_fail(c, APM_E_STATE, "stat records are not retained (APM_F_KEEP_STATS)");
    apm_status st = fetch_dense(c, c->d_stats, c->K, out, cap, n);
    if (!st) c->counters.stat_records += *n;
    return st;
}

apm_status apm_fetch_fs(apm_ctx* c, apm_fs_record* out, uint64_t cap, uint64_t* n) {
    if (!c || !n) return APM_E_INVALID;
    APM_ENTER(c);
    if (!(c->cfg.flags & APM_F_KEEP_FS) || !with_z(c)) return apm_fail(c, APM_E_STATE, "fs records are not retained (APM_F_KEEP_FS | APM_F_ZSCORE)");
    const size_t per_tick = (size_t)c->K * c->cfg.num_lags, total = (size_t)c->ticks_buffered * per_tick;
    if (out && out == c->h_fs_bound && total <= c->h_fs_bound_cap) {
        // the records were streamed here tick by tick: wait for the last copy and drop invalid ones in place
        *n = 0;
        apm_status st = sync_all(c);
        if (st) return st;
        if (cap < total) return apm_fail(c, APM_E_CAPACITY, "output buffer smaller than ticks_buffered * records per tick");
        uint32_t unseen = 1;
        CUDA_TRY(c, cudaMemcpy(&unseen, &c->d_counters->unseen_records, sizeof unseen, cudaMemcpyDeviceToHost));
        size_t w = total;
        if (unseen) {  // some keys have not been seen yet: their slots hold invalid records
            w = 0;
            for (size_t i = 0; i < total; ++i)
                if (out[i].flags & 1) {
                    if (w != i) out[w] = out[i];
                    ++w;
                }
        }
        *n = w;
        c->counters.fs_records += w;
        return APM_OK;
    }
    apm_status st = fetch_dense(c, c->d_fs, per_tick, out, cap, n);
    if (!st) c->counters.fs_records += *n;
    return st;
}

apm_status apm_bind_fs_output(apm_ctx* c, apm_fs_record* pinned, uint64_t cap_records) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    apm_status st = sync_all(c);
    if (st) return st;
    if (pinned) {
        if (!(c->cfg.flags & APM_F_KEEP_FS) || !with_z(c)) return apm_fail(c, APM_E_STATE, "bind_fs_output needs APM_F_KEEP_FS | APM_F_ZSCORE");
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, pinned) != cudaSuccess || at.type != cudaMemoryTypeHost) {
            cudaGetLastError();
            return apm_fail(c, APM_E_INVALID, "bind_fs_output: the buffer must be page-locked host memory (apm_host_alloc)");
        }
    }
    c->h_fs_bound = pinned;
    c->h_fs_bound_cap = pinned ? cap_records : 0;
    return APM_OK;
}

apm_status apm_fetch_alerts(apm_ctx* c, apm_fs_record* out, uint64_t cap, uint64_t* n) {
    if (!c || !n) return APM_E_INVALID;
    APM_ENTER(c);
    *n = 0;
    if (!c->d_alerts) return apm_fail(c, APM_E_STATE, "alerts are not compacted (APM_F_ALERTS | APM_F_ZSCORE)");
    if (c->ticks_buffered == 0) return APM_OK;
    std::vector<unsigned int> counts(c->ticks_buffered);
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));  // the alert blocks are written on the baseline stream
    CUDA_TRY(c, cudaMemcpyAsync(counts.data(), c->d_alert_count, sizeof(unsigned int) * c->ticks_buffered, cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    uint64_t total = 0;
    for (unsigned int k : counts) {
        if (k > c->cfg.max_alerts_per_tick) return apm_fail(c, APM_E_CAPACITY, "more alerts in one tick than max_alerts_per_tick");
        total += k;
    }
    if (total > cap) return apm_fail(c, APM_E_CAPACITY, "alert output buffer too small");
    uint64_t w = 0;
    for (uint32_t t = 0; t < c->ticks_buffered; ++t) {
        if (!counts[t]) continue;
        // each tick's block is already (lag, key)-ascending (ordered compaction)
        CUDA_TRY(c, cudaMemcpyAsync(out + w, c->d_alerts + (size_t)t * c->cfg.max_alerts_per_tick, counts[t] * sizeof(apm_fs_record), cudaMemcpyDeviceToHost, c->s_main));
        w += counts[t];
    }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    c->counters.d2h_bytes += total * sizeof(apm_fs_record) + counts.size() * 4;
    c->counters.alerts += total;
    *n = total;
    return APM_OK;
}

apm_status apm_release_ticks(apm_ctx* c) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (c->d_alert_count) {
        { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
        CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
        { apm_status cr_ = apm_comm_release(c); if (cr_) return cr_; }  // the gather stream still reads the alert blocks
        CUDA_TRY(c, cudaMemsetAsync(c->d_alert_count, 0, sizeof(unsigned int) * c->cfg.out_ticks, c->s_main));
        CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    }
    CUDA_TRY(c, cudaMemsetAsync(&c->d_counters->unseen_records, 0, sizeof(uint32_t), c->s_main));
    c->ticks_buffered = 0;
    c->pt_out_used = 0;
    return APM_OK;
}

apm_status apm_ring_load(apm_ctx* c, uint32_t key, int32_t l, int32_t series, const double* values, uint32_t len) {
    if (!c || (!values && len)) return APM_E_INVALID;
    APM_ENTER(c);
    if (!with_z(c) || l < 0 || l >= c->cfg.num_lags || series < 0 || series >= APM_NUM_SERIES || key >= c->K) return apm_fail(c, APM_E_INVALID, "ring_load: index");
    const uint32_t L = (uint32_t)c->cfg.lags[l];
    if (len > L) return apm_fail(c, APM_E_INVALID, "ring_load: list longer than its lag");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    const size_t ncol = c->ring_pitch, col = (size_t)series * c->K + key;
    const uint32_t head = c->ring_head[l];
    // the list ends just before `head`: element i (oldest first) lives at slot head - len + i
    std::vector<double> tmp(values, values + len);
    const uint32_t first = (head + L - len) % L;
    const uint32_t n1 = std::min(len, L - first);
    if (n1) CUDA_TRY(c, cudaMemcpy2D(c->ring[l] + (size_t)first * ncol + col, ncol * sizeof(double), tmp.data(), sizeof(double), sizeof(double), n1, cudaMemcpyHostToDevice));
    if (len > n1) CUDA_TRY(c, cudaMemcpy2D(c->ring[l] + col, ncol * sizeof(double), tmp.data() + n1, sizeof(double), sizeof(double), len - n1, cudaMemcpyHostToDevice));
    CUDA_TRY(c, cudaMemcpy(c->ring_len[l] + col, &len, sizeof(uint32_t), cudaMemcpyHostToDevice));
    CUDA_TRY(c, cudaMemcpy(c->ring_next[l] + col, &head, sizeof(uint32_t), cudaMemcpyHostToDevice));  // the list ends just before `head`
    c->may_misalign = true;
    c->misalign_mark = c->tick_seq;
    c->inc_dirty[l] = true;
    return APM_OK;
}

apm_status apm_ring_load_all(apm_ctx* c, int32_t l, const double* values, uint32_t len, int32_t on_device) {
    if (!c || (!values && len)) return APM_E_INVALID;
    APM_ENTER(c);
    if (!with_z(c) || l < 0 || l >= c->cfg.num_lags) return apm_fail(c, APM_E_INVALID, "ring_load_all: index");
    const uint32_t L = (uint32_t)c->cfg.lags[l];
    if (len > L) return apm_fail(c, APM_E_INVALID, "ring_load_all: lists longer than the lag");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    const size_t ncol = 3 * (size_t)c->K, row = ncol * sizeof(double), pitch = c->ring_pitch * sizeof(double);
    const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    const uint32_t head = c->ring_head[l], first = (head + L - len) % L, n1 = std::min(len, L - first);
    if (n1) CUDA_TRY(c, cudaMemcpy2D(c->ring[l] + (size_t)first * c->ring_pitch, pitch, values, row, row, n1, kind));
    if (len > n1) CUDA_TRY(c, cudaMemcpy2D(c->ring[l], pitch, values + (size_t)n1 * ncol, row, row, len - n1, kind));
    std::vector<uint32_t> lens(ncol, len);
    CUDA_TRY(c, cudaMemcpy(c->ring_len[l], lens.data(), ncol * sizeof(uint32_t), cudaMemcpyHostToDevice));
    std::vector<uint32_t> nexts(ncol, head);
    CUDA_TRY(c, cudaMemcpy(c->ring_next[l], nexts.data(), ncol * sizeof(uint32_t), cudaMemcpyHostToDevice));
    c->may_misalign = true;
    c->misalign_mark = c->tick_seq;
    c->inc_dirty[l] = true;
    return APM_OK;
}

apm_status apm_ring_read(apm_ctx* c, uint32_t key, int32_t l, int32_t series, double* values, uint32_t cap, uint32_t* len_out) {
    if (!c || !len_out) return APM_E_INVALID;
    APM_ENTER(c);
    if (!with_z(c) || l < 0 || l >= c->cfg.num_lags || series < 0 || series >= APM_NUM_SERIES || key >= c->K) return apm_fail(c, APM_E_INVALID, "ring_read: index");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    const uint32_t L = (uint32_t)c->cfg.lags[l];
    const size_t ncol = c->ring_pitch, col = (size_t)series * c->K + key;
    uint32_t len = 0;
    CUDA_TRY(c, cudaMemcpy(&len, c->ring_len[l] + col, sizeof(uint32_t), cudaMemcpyDeviceToHost));
    *len_out = len;
    if (len == 0) return APM_OK;
    if (!values || cap < len) return apm_fail(c, APM_E_CAPACITY, "ring_read: buffer too small");
    uint32_t head = 0;  // the slot this column's next push belongs in: a silent key's list does not follow the global head
    CUDA_TRY(c, cudaMemcpy(&head, c->ring_next[l] + col, sizeof(uint32_t), cudaMemcpyDeviceToHost));
    const uint32_t first = (head + L - len) % L;
    const uint32_t n1 = std::min(len, L - first);
    CUDA_TRY(c, cudaMemcpy2D(values, sizeof(double), c->ring[l] + (size_t)first * ncol + col, ncol * sizeof(double), sizeof(double), n1, cudaMemcpyDeviceToHost));
    if (len > n1) CUDA_TRY(c, cudaMemcpy2D(values + n1, sizeof(double), c->ring[l] + col, ncol * sizeof(double), sizeof(double), len - n1, cudaMemcpyDeviceToHost));
    return APM_OK;
}

int64_t apm_latest(const apm_ctx* c) { return c ? c->latest : 0; }

apm_status apm_set_latest(apm_ctx* c, int64_t latest) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (latest <= 0) return apm_fail(c, APM_E_INVALID, "set_latest: label must be positive");
    apm_status st = sync_all(c);
    if (st) return st;
    c->latest = latest;
    if (c->pt_cap && c->pt_base_ms == INT64_MIN) c->pt_base_ms = latest * c->cfg.bucket_ms - (1ll << 39);  // as the first tick would
    uint64_t mask = 0;
    for (int s = 0; s < c->S; ++s) c->slot_label[s] = -1;
    for (int64_t lab = std::max<int64_t>(0, latest - c->keep); lab <= latest; ++lab) {
        const int slot = (int)(lab % c->S);
        c->slot_label[slot] = lab;
        mask |= 1ull << slot;
        c->host_count[slot] = c->host_sorted[slot] = 0;
    }
    st = apm_launch_clear_slots(c, mask);
    if (st) return st;
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    return APM_OK;
}

apm_status apm_mark_seen(apm_ctx* c, uint32_t key_lo, uint32_t n) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if ((uint64_t)key_lo + n > c->K) return apm_fail(c, APM_E_INVALID, "mark_seen: range");
    apm_status st = sync_all(c);
    if (st) return st;
    if (n) CUDA_TRY(c, cudaMemset(c->seen + key_lo, 1, n));
    return APM_OK;
}

apm_status apm_log_read(apm_ctx* c, int64_t label, uint64_t* out, uint64_t cap, uint64_t* n) {
    if (!c || !n) return APM_E_INVALID;
    APM_ENTER(c);
    *n = 0;
    if (!c->log_cap) return apm_fail(c, APM_E_STATE, "value logs are only kept with APM_F_PERCENTILES");
    apm_status st = sync_all(c);
    if (st) return st;
    if (label < 0) return APM_OK;
    const int slot = (int)(label % c->S);
    if (c->slot_label[slot] != label) return APM_OK;  // not retained
    SlotMeta m;
    CUDA_TRY(c, cudaMemcpy(&m, c->slot_meta + slot, sizeof m, cudaMemcpyDeviceToHost));
    const uint64_t cnt = std::min<uint64_t>(m.count, c->log_cap);
    *n = cnt;
    if (cnt == 0 || !out) return APM_OK;  // out == NULL: size query
    if (cap < cnt) return apm_fail(c, APM_E_CAPACITY, "log_read: buffer too small");
    CUDA_TRY(c, cudaMemcpy(out, c->log_buf[slot], cnt * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return APM_OK;
}

apm_status apm_get_counters(apm_ctx* c, apm_counters* out) {
    if (!c || !out) return APM_E_INVALID;
    APM_ENTER(c);
    apm_status st = refresh_counters(c);
    // keys_seen: count on the host (called rarely)
    std::vector<uint8_t> seen(c->K);
    CUDA_TRY(c, cudaMemcpy(seen.data(), c->seen, c->K, cudaMemcpyDeviceToHost));
    uint64_t ks = 0;
    for (uint8_t s : seen) ks += s;
    c->counters.keys_seen = ks;
    *out = c->counters;
    return st;
}

apm_status apm_reset_counters(apm_ctx* c) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    CUDA_TRY(c, cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->s_main));
    memset(&c->counters, 0, sizeof c->counters);
    c->launches = 0;
    return APM_OK;
}

uint32_t apm_owner_of_key(uint64_t k, uint32_t nranks) {
    // splitmix64 finaliser
    k += 0x9E3779B97F4A7C15ull;
    k = (k ^ (k >> 30)) * 0xBF58476D1CE4E5B9ull;
    k = (k ^ (k >> 27)) * 0x94D049BB133111EBull;
    k ^= k >> 31;
    return nranks ? (uint32_t)(k % nranks) : 0;
}

// Count-aware placement (host arithmetic only; the same algorithm as sharding.balanced_owner_table): keys that alone
// exceed 1 / (8 nranks) of the load go longest-first to the least loaded rank, the light ones keep their hash owner
// unless that rank is already above the mean.  load(key) = count + key_cost.
apm_status apm_owner_table_build(const uint64_t* key_counts, uint32_t num_keys, uint32_t nranks, uint64_t key_cost,
                                 uint32_t* owner_out) {
    if ((!key_counts || !owner_out) && num_keys) return apm_fail(nullptr, APM_E_INVALID, "apm_owner_table_build: null argument");
    const uint32_t n = nranks ? nranks : 1;
    double total = 0.0;
    std::vector<double> w(num_keys);
    for (uint32_t k = 0; k < num_keys; ++k) {
        w[k] = (double)key_counts[k] + (double)key_cost;
        total += w[k];
        owner_out[k] = apm_owner_of_key(k, n);
    }
    if (n == 1 || !(total > 0.0)) return APM_OK;
    std::vector<double> load(n, 0.0);
    auto least = [&]() { return (uint32_t)(std::min_element(load.begin(), load.end()) - load.begin()); };
    const double heavy_above = total / (8.0 * n), mean = total / n;
    std::vector<uint32_t> heavy, light;
    for (uint32_t k = 0; k < num_keys; ++k) (w[k] > heavy_above ? heavy : light).push_back(k);
    auto by_load = [&](uint32_t a, uint32_t b) { return w[a] > w[b]; };
    std::stable_sort(heavy.begin(), heavy.end(), by_load);
    std::stable_sort(light.begin(), light.end(), by_load);
    for (uint32_t k : heavy) {
        const uint32_t r = least();
        owner_out[k] = r;
        load[r] += w[k];
    }
    for (uint32_t k : light) {
        uint32_t r = owner_out[k];
        if (load[r] + w[k] > mean) {
            r = least();
            owner_out[k] = r;
        }
        load[r] += w[k];
    }
    return APM_OK;
}

// Placement that evens out the TRANSACTIONS while every rank keeps (almost) the same number of KEYS: longest-first onto
// the rank with the fewest transactions among those that still have room for a key.  For Zipf traffic over many keys -
// one key can be most of a rank's fair share - where a rank's memory and its baseline walk follow the key count and its
// ingest / sort / PCIe time the transactions.
apm_status apm_owner_table_build_capped(const uint64_t* key_counts, uint32_t num_keys, uint32_t nranks, double key_slack,
                                        uint32_t* owner_out) {
    if ((!key_counts || !owner_out) && num_keys) return apm_fail(nullptr, APM_E_INVALID, "apm_owner_table_build_capped: null argument");
    const uint32_t n = nranks ? nranks : 1;
    if (!(key_slack >= 0.0)) key_slack = 0.0;
    const uint64_t cap = (uint64_t)std::ceil((double)num_keys / n * (1.0 + key_slack));
    std::vector<uint32_t> order(num_keys);
    for (uint32_t k = 0; k < num_keys; ++k) order[k] = k;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return key_counts[a] > key_counts[b]; });
    std::vector<uint64_t> load(n, 0), held(n, 0);
    for (uint32_t k : order) {
        uint32_t best = n;
        for (uint32_t r = 0; r < n; ++r)
            if (held[r] < cap && (best == n || load[r] < load[best])) best = r;
        if (best == n) best = 0;  // cannot happen: n * cap >= num_keys
        owner_out[k] = best;
        load[best] += key_counts[k];
        held[best] += 1;
    }
    return APM_OK;
}

// Host binning for key-sharded deployments (pure host arithmetic).  One pass over the arrival-ordered global stream:
// every valid transaction goes to the rank that owns its key, arrival order kept inside a rank, and every rank learns
// where the GLOBAL stream ticks - the running maximum of the bucket label over all transactions, whoever owns them
// (stream_calc_stats.js:348) - as the number of its own transactions that precede the trigger.
apm_status apm_partition_host(const apm_tx_columns* cols, uint64_t n, int64_t bucket_ms, int64_t latest_bucket,
                              uint32_t num_services, uint32_t num_jvms, uint32_t nranks, const uint32_t* owner_table,
                              const uint32_t* local_id, apm_partition_out* out, int64_t* latest_out) {
    if ((!cols && n) || !out || nranks == 0 || bucket_ms <= 0 || num_services == 0 || num_jvms == 0)
        return apm_fail(nullptr, APM_E_INVALID, "apm_partition_host: bad argument");
    for (uint32_t r = 0; r < nranks; ++r) { out[r].n = 0; out[r].num_marks = 0; }
    int64_t latest = latest_bucket;
    for (uint64_t i = 0; i < n; ++i) {
        const uint32_t sv = cols->service_id[i], jv = cols->jvm_id[i];
        const int64_t ts = cols->end_ts_ms[i];
        const int64_t lab = ts >= 0 ? ts / bucket_ms : -1;
        if (lab > latest) {  // the tick fires before this transaction is added, on every rank
            latest = lab;
            for (uint32_t r = 0; r < nranks; ++r) {
                apm_partition_out& o = out[r];
                if (o.num_marks >= o.marks_cap) return apm_fail(nullptr, APM_E_CAPACITY, "apm_partition_host: marks_cap too small");
                o.marks[o.num_marks].pos = o.n;
                o.marks[o.num_marks].label = lab;
                o.num_marks++;
            }
        }
        uint32_t r = 0, s_out = sv, j_out = jv;
        if (sv < num_services && jv < num_jvms) {
            const uint64_t key = (uint64_t)jv * num_services + sv;
            r = owner_table ? owner_table[key] : apm_owner_of_key(key, nranks);
            if (r >= nranks) return apm_fail(nullptr, APM_E_INVALID, "apm_partition_host: owner_table entry out of range");
            if (local_id) { s_out = local_id[key]; j_out = 0; }
        }  // else: an invalid transaction; rank 0 counts it
        apm_partition_out& o = out[r];
        if (o.n >= o.cap) return apm_fail(nullptr, APM_E_CAPACITY, "apm_partition_host: a rank's output columns are full");
        o.cols.service_id[o.n] = s_out;
        o.cols.jvm_id[o.n] = j_out;
        o.cols.end_ts_ms[o.n] = ts;
        o.cols.elapsed_ms[o.n] = cols->elapsed_ms[i];
        o.n++;
    }
    if (latest_out) *latest_out = latest;
    return APM_OK;
}

}  // extern "C"

// Device-side stopwatch on the context's stream (what bench.py brackets its timed region with).
extern "C" apm_status apm_timer_start(apm_ctx* c) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (!c->timer[0]) { CUDA_TRY(c, cudaEventCreate(&c->timer[0])); CUDA_TRY(c, cudaEventCreate(&c->timer[1])); }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_copy));
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaEventRecord(c->timer[0], c->s_main));
    return APM_OK;
}

extern "C" apm_status apm_timer_stop(apm_ctx* c, double* ms) {
    if (!c || !ms) return APM_E_INVALID;
    APM_ENTER(c);
    if (!c->timer[0]) return apm_fail(c, APM_E_STATE, "timer_stop without timer_start");
    CUDA_TRY(c, cudaStreamSynchronize(c->s_copy));
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));  // the stop event on s_main then follows all work of both streams
    CUDA_TRY(c, cudaEventRecord(c->timer[1], c->s_main));
    CUDA_TRY(c, cudaEventSynchronize(c->timer[1]));
    float f = 0;
    CUDA_TRY(c, cudaEventElapsedTime(&f, c->timer[0], c->timer[1]));
    *ms = f;
    return APM_OK;
}

extern "C" apm_status apm_profile_enable(apm_ctx* c, int32_t on) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    c->prof_on = on != 0;
    return APM_OK;
}

extern "C" apm_status apm_profile_read(apm_ctx* c, apm_kernel_times* out) {
    if (!c || !out) return APM_E_INVALID;
    APM_ENTER(c);
    memset(out, 0, sizeof *out);
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    // APM_TIMELINE_FILE=<path>: every span as "group start_us end_us" relative to the first span of this read (where the
    // streams overlap, where they wait for each other)
    if (const char* path = getenv("APM_TIMELINE_FILE")) {
        if (FILE* f = c->prof_spans.empty() ? nullptr : fopen(path, "a")) {
            const cudaEvent_t base = c->prof_spans.front().a;
            for (auto& sp : c->prof_spans) {
                float a_ms = 0, b_ms = 0;
                if (cudaEventElapsedTime(&a_ms, base, sp.a) == cudaSuccess && cudaEventElapsedTime(&b_ms, base, sp.b) == cudaSuccess)
                    fprintf(f, "%d %.2f %.2f\n", sp.group, a_ms * 1e3, b_ms * 1e3);
            }
            fprintf(f, "-1 0 0\n");
            fclose(f);
        }
    }
    for (auto& sp : c->prof_spans) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess && sp.group < APM_KT_GROUPS) {
            out->ms[sp.group] += ms;
            out->spans[sp.group] += 1;
        }
        c->prof_pool.push_back(sp.a);
        c->prof_pool.push_back(sp.b);
    }
    c->prof_spans.clear();
    return APM_OK;
}

// Run the rolling baseline on its own stream under the next tick's ingest/sort (default), or in line.
extern "C" apm_status apm_set_overlap(apm_ctx* c, int32_t on) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    apm_status st = sync_all(c);
    if (st) return st;
    c->overlap = on != 0;
    return APM_OK;
}

// Page-locked host buffers for results: a fetch into one of these is a plain DMA.
extern "C" apm_status apm_host_alloc(uint64_t bytes, void** out) {
    if (!out) return APM_E_INVALID;
    *out = nullptr;
    cudaError_t e = cudaMallocHost(out, bytes ? bytes : 1);
    if (e != cudaSuccess) return apm_fail_cuda(nullptr, e, "cudaMallocHost", __FILE__, __LINE__);
    return APM_OK;
}

extern "C" void apm_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

extern "C" uint32_t apm_ticks_buffered(const apm_ctx* c) { return c ? c->ticks_buffered : 0; }

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_capi() {
    const void* fns[] = {(const void*)widen_tx8_kernel, (const void*)widen_tx_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
