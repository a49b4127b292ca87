// Checkpoint / resume: replaces saveToResumeFile / resumeFromFileIfExists
// (stream_calc_stats.js:54-87: servers{}.services{}.buckets{} + latestBucket;
//  stream_calc_z_score.js:37-64: servers{}.services{}.lags{}.{avg,per75,per95}List)
// with a binary image of the same state: bucket accumulators, value logs, seen keys,
// lag rings, list lengths, per-key settings and - with APM_F_PASSTHROUGH - the pending words of the
// device-side reorder buffer (the reference's minHeap); without that flag the heap stays with the host.
#include <cstdio>
#include <cstring>
#include <vector>

#include "ctx.h"

namespace {

struct StateHeader {
    char magic[8];  // "APMB200\0"
    uint32_t version;
    uint32_t K;
    int32_t S;
    int32_t num_lags;
    int64_t bucket_ms;
    int32_t window, buffer;
    int32_t lags[APM_MAX_LAGS];
    uint32_t ring_head[APM_MAX_LAGS];
    int64_t latest;
    int64_t slot_label[APM_MAX_SLOTS];
    uint64_t log_count[APM_MAX_SLOTS];
    uint32_t has_logs, has_rings;
    uint32_t max_value, pad_;  // largest duration in any value log (bounds the radix passes after a resume)
    // the device-side reorder buffer of the transaction passthrough (APM_F_PASSTHROUGH): what the reference keeps as
    // `minHeap` in the stats resume file and re-pushes on resume (stream_calc_stats.js:40-52, 54-61)
    uint32_t has_pt, pad2_;
    uint64_t pt_count;         // pending words
    uint64_t pt_seq;           // next arrival sequence number
    int64_t pt_base_ms;        // time base of the pending words
};

struct File {
    FILE* f = nullptr;
    ~File() { if (f) fclose(f); }
};

bool put(FILE* f, const void* dev, size_t bytes, std::vector<char>& tmp) {
    if (!bytes) return true;
    tmp.resize(bytes);
    if (cudaMemcpy(tmp.data(), dev, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) return false;
    return fwrite(tmp.data(), 1, bytes, f) == bytes;
}

bool get(FILE* f, void* dev, size_t bytes, std::vector<char>& tmp) {
    if (!bytes) return true;
    tmp.resize(bytes);
    if (fread(tmp.data(), 1, bytes, f) != bytes) return false;
    return cudaMemcpy(dev, tmp.data(), bytes, cudaMemcpyHostToDevice) == cudaSuccess;
}

}  // namespace

extern "C" {

apm_status apm_state_save(apm_ctx* c, const char* path) {
    if (!c || !path) return APM_E_INVALID;
    APM_ENTER(c);
    CUDA_TRY(c, cudaStreamSynchronize(c->s_copy));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    File fl;
    fl.f = fopen(path, "wb");
    if (!fl.f) return apm_fail(c, APM_E_IO, std::string("cannot open for writing: ") + path);
    StateHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, "APMB200", 8);
    h.version = 6; h.K = c->K; h.S = c->S; h.num_lags = c->log_cap ? c->cfg.num_lags : c->cfg.num_lags;
    h.bucket_ms = c->cfg.bucket_ms; h.window = c->cfg.window_buckets; h.buffer = c->cfg.buffer_buckets;
    h.latest = c->latest;
    const bool rings = (c->cfg.flags & APM_F_ZSCORE) && c->cfg.num_lags > 0;
    h.has_logs = c->log_cap ? 1 : 0;
    h.has_rings = rings ? 1 : 0;
    h.max_value = c->known_max_value;
    for (int l = 0; l < c->cfg.num_lags; ++l) { h.lags[l] = c->cfg.lags[l]; h.ring_head[l] = c->ring_head[l]; }
    std::vector<SlotMeta> meta(c->S);
    CUDA_TRY(c, cudaMemcpy(meta.data(), c->slot_meta, sizeof(SlotMeta) * c->S, cudaMemcpyDeviceToHost));
    for (int s = 0; s < c->S; ++s) { h.slot_label[s] = c->slot_label[s]; h.log_count[s] = std::min<uint64_t>(meta[s].count, c->log_cap); }
    if (c->pt_cap) {
        unsigned long long cur = 0;
        CUDA_TRY(c, cudaMemcpy(&cur, c->pt_cursors, sizeof cur, cudaMemcpyDeviceToHost));
        if (cur > c->pt_cap) return apm_fail(c, APM_E_CAPACITY, "passthrough: more pending transactions than the reorder buffer holds");
        h.has_pt = 1;
        h.pt_count = cur;
        h.pt_seq = c->pt_seq;
        h.pt_base_ms = c->pt_base_ms;
    }
    std::vector<char> tmp;
    bool ok = fwrite(&h, sizeof h, 1, fl.f) == 1;
    ok = ok && put(fl.f, c->acc_raw, (size_t)APM_ACC_BYTES_PER_ENTRY * c->S * c->K, tmp);
    ok = ok && put(fl.f, c->seen, c->K, tmp);
    ok = ok && put(fl.f, c->slot_meta, sizeof(SlotMeta) * c->S, tmp);
    if (h.has_logs)
        for (int s = 0; s < c->S && ok; ++s) ok = put(fl.f, c->log_buf[s], h.log_count[s] * sizeof(uint64_t), tmp);
    if (rings)
        for (int l = 0; l < c->cfg.num_lags && ok; ++l) {
            ok = put(fl.f, c->ring[l], sizeof(double) * (size_t)c->cfg.lags[l] * c->ring_pitch, tmp);
            ok = ok && put(fl.f, c->ring_len[l], sizeof(uint32_t) * 3 * c->K, tmp);
            ok = ok && put(fl.f, c->ring_next[l], sizeof(uint32_t) * 3 * c->K, tmp);
            ok = ok && put(fl.f, c->thr[l], sizeof(double) * c->K, tmp);
            ok = ok && put(fl.f, c->infl[l], sizeof(double) * c->K, tmp);
        }
    if (rings) ok = ok && put(fl.f, c->alert_count, sizeof(uint32_t) * c->cfg.num_lags * c->K, tmp);
    if (h.has_pt) ok = ok && put(fl.f, c->pt_pend[c->pt_cur], h.pt_count * sizeof(uint64_t), tmp);
    if (!ok) return apm_fail(c, APM_E_IO, std::string("short write: ") + path);
    return APM_OK;
}

apm_status apm_state_load(apm_ctx* c, const char* path) {
    if (!c || !path) return APM_E_INVALID;
    APM_ENTER(c);
    CUDA_TRY(c, cudaStreamSynchronize(c->s_copy));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    File fl;
    fl.f = fopen(path, "rb");
    if (!fl.f) return apm_fail(c, APM_E_IO, std::string("cannot open for reading: ") + path);
    StateHeader h;
    if (fread(&h, sizeof h, 1, fl.f) != 1 || memcmp(h.magic, "APMB200", 8) != 0 || h.version != 6)
        return apm_fail(c, APM_E_IO, "not an apm_b200 state file");
    const bool rings = (c->cfg.flags & APM_F_ZSCORE) && c->cfg.num_lags > 0;
    bool same = h.K == c->K && h.S == c->S && h.bucket_ms == c->cfg.bucket_ms && h.window == c->cfg.window_buckets &&
                h.buffer == c->cfg.buffer_buckets && h.num_lags == c->cfg.num_lags && (h.has_logs != 0) == (c->log_cap != 0) &&
                (h.has_rings != 0) == rings && (h.has_pt != 0) == (c->pt_cap != 0);
    for (int l = 0; same && l < c->cfg.num_lags; ++l) same = h.lags[l] == c->cfg.lags[l];
    if (!same) return apm_fail(c, APM_E_INVALID, "state file was written with a different configuration");
    for (int s = 0; s < c->S; ++s)
        if (h.log_count[s] > c->log_cap) return apm_fail(c, APM_E_CAPACITY, "state file holds a value log above max_tx_per_bucket");
    if (h.has_pt && h.pt_count > c->pt_cap) return apm_fail(c, APM_E_CAPACITY, "state file holds more pending passthrough transactions than the reorder buffer");
    std::vector<char> tmp;
    bool ok = get(fl.f, c->acc_raw, (size_t)APM_ACC_BYTES_PER_ENTRY * c->S * c->K, tmp);
    ok = ok && get(fl.f, c->seen, c->K, tmp);
    ok = ok && get(fl.f, c->slot_meta, sizeof(SlotMeta) * c->S, tmp);
    if (h.has_logs)
        for (int s = 0; s < c->S && ok; ++s) ok = get(fl.f, c->log_buf[s], h.log_count[s] * sizeof(uint64_t), tmp);
    if (rings)
        for (int l = 0; l < c->cfg.num_lags && ok; ++l) {
            ok = get(fl.f, c->ring[l], sizeof(double) * (size_t)c->cfg.lags[l] * c->ring_pitch, tmp);
            ok = ok && get(fl.f, c->ring_len[l], sizeof(uint32_t) * 3 * c->K, tmp);
            ok = ok && get(fl.f, c->ring_next[l], sizeof(uint32_t) * 3 * c->K, tmp);
            ok = ok && get(fl.f, c->thr[l], sizeof(double) * c->K, tmp);
            ok = ok && get(fl.f, c->infl[l], sizeof(double) * c->K, tmp);
        }
    if (rings) ok = ok && get(fl.f, c->alert_count, sizeof(uint32_t) * c->cfg.num_lags * c->K, tmp);
    if (h.has_pt) {
        ok = ok && get(fl.f, c->pt_pend[c->pt_cur], h.pt_count * sizeof(uint64_t), tmp);
        const unsigned long long cur[3] = {h.pt_count, 0ull, 0ull};
        ok = ok && cudaMemcpy(c->pt_cursors, cur, sizeof cur, cudaMemcpyHostToDevice) == cudaSuccess;
        c->pt_seq = h.pt_seq;
        c->pt_base_ms = h.pt_base_ms;
        c->pt_out_used = 0;
    }
    if (!ok) return apm_fail(c, APM_E_IO, std::string("short read: ") + path);
    c->latest = h.latest;
    // a context saved before its first tick has no time base yet: give it the one the first tick would have chosen
    if (c->pt_cap && c->pt_base_ms == INT64_MIN && h.latest > 0) c->pt_base_ms = h.latest * c->cfg.bucket_ms - (1ll << 39);
    for (int s = 0; s < c->S; ++s) c->slot_label[s] = h.slot_label[s];
    {   // the host's view of the value logs
        std::vector<SlotMeta> meta(c->S);
        CUDA_TRY(c, cudaMemcpy(meta.data(), c->slot_meta, sizeof(SlotMeta) * c->S, cudaMemcpyDeviceToHost));
        for (int s = 0; s < c->S; ++s) { c->host_count[s] = meta[s].count; c->host_sorted[s] = meta[s].sorted; }
        c->known_max_value = h.max_value;
    }
    for (int l = 0; l < c->cfg.num_lags; ++l) c->ring_head[l] = h.ring_head[l];
    c->ticks_buffered = 0;
    for (int l = 0; l < APM_MAX_LAGS; ++l) c->inc_dirty[l] = true;
    c->may_misalign = true;  // loaded lists of keys that are silent now re-enter the lock step when they emit
    c->misalign_mark = c->tick_seq;
    return APM_OK;
}

}  // extern "C"
