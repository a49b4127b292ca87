// Context layout and the host-side launchers of each kernel group.
#pragma once
#include <climits>
#include <cstdint>
#include <string>
#include <vector>

#include <cuda.h>

#include "common.cuh"

// k3_groupsort.cu: which of the two plan buffers is current and how many splitters it holds (0: no plan)
#define APM_GS_MAX_SPLIT 2048
struct apm_groupsort_state {
    int cur = 0;
    uint32_t P = 0;
};

struct TickMark {
    unsigned long long pos;  // index in the chunk of the transaction that triggers the tick
    long long label;         // its bucket label (the new latestBucket)
};

struct PendingBaseline {
    bool valid;
    int64_t tick_ts;
    uint32_t fslot, tslot;
    int buf, ev;
};

struct apm_ctx {
    apm_config cfg;
    uint32_t K = 0;       // keys = num_services * num_jvms
    int num_sms = 148;    // SMs of cfg.device
    int S = 0;            // bucket slots = window + buffer + 1 (labels latest-keep .. latest)
    int keep = 0;         // NUM_KEEP_INTERVALS = window + buffer (stream_calc_stats.js:233)
    int64_t latest = 0;   // latestBucket (:42)
    double inv_bucket = 0, tpm_div = 1;
    double pfrac[APM_MAX_PCT] = {0, 0, 0, 0};  // percentile / 100.0 (util_methods.js:129)

    // s_main: ingest, sort, window statistics.  s_base: rolling baseline + alert compaction of tick i,
    // which overlaps the ingest/sort of tick i+1 (linked by ev_stats / ev_base).  s_copy: H2D staging.
    cudaStream_t s_main = nullptr, s_copy = nullptr, s_base = nullptr;
    cudaEvent_t ev_stats[4] = {nullptr, nullptr, nullptr, nullptr}, ev_base[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_mid = nullptr;
    PendingBaseline pending = {false, 0, 0, 0, 0, 0};  // baseline of the last tick, not launched yet
    uint64_t tick_seq = 0;        // ticks run so far (selects the zin / emit buffer and the events)
    bool overlap = true;          // APM_NO_OVERLAP=1 runs the baseline on s_main (A/B measurements)
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};

    // --- bucket accumulators and value logs
    void* acc_raw = nullptr;     // one allocation behind the six accumulator planes
    AccSoA acc = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // each [S][K]
    uint8_t* seen = nullptr;     // [K] key has ever been ingested
    SlotMeta* slot_meta = nullptr;  // [S]
    int64_t slot_label[APM_MAX_SLOTS];
    uint64_t* log_buf[APM_MAX_SLOTS + 1];  // S value logs + 1 spare for the sort ping-pong
    uint64_t** d_log_ptrs = nullptr;       // device copy of log_buf[0..S)
    uint64_t log_cap = 0;
    // radix sort scratch
    uint32_t* sort_hist = nullptr;  // [256][tiles]
    uint32_t* sort_base = nullptr;  // [256]
    uint64_t sort_tiles_cap = 0;
    unsigned char* sort_scratch = nullptr;  // onesweep: tickets, global histograms, per-tile look-back status
    size_t sort_scratch_bytes = 0;
    bool sort_legacy = false;               // APM_SORT_LEGACY=1: three-kernel-per-pass sort (A/B measurements)
    bool small_tick = true;                 // APM_SMALL_TICK=0: separate clear / sort / rebuild launches also for small shapes
    bool sort_group = false;                // APM_SORT_GROUP=1: planned partition + shared-memory group sorts (k3_groupsort.cu;
                                            // measured no faster than the radix passes, kept as an A/B path)
    apm_groupsort_state gs;                 // its plan, built from the last bucket that was sorted
    unsigned char* gs_scratch = nullptr;

    // --- per-chunk pre-pass (tick finder + segment table).  Two sets: while the ticks of chunk i are being queued,
    //     the pre-pass of chunk i + 1 runs on its own stream (s_pre), so a chunk boundary costs the ingest / sort
    //     chain neither the pre-pass kernels nor the host round trips.
    struct PrePass {
        long long* tf_tile_max = nullptr;
        long long* tf_tile_prefix = nullptr;
        TickMark* d_ticks = nullptr;
        unsigned int* d_nticks = nullptr;
        TickMark* h_ticks = nullptr;        // pinned; sorted by position once the pre-pass has finished
        unsigned int* h_nticks = nullptr;   // pinned
        // log entries per (segment, bucket age): from it the host knows the exact length of every bucket's value
        // log, so the tick loop never waits for the device
        uint32_t* d_seg_counts = nullptr;   // [(max_ticks_per_chunk + 1)][keep + 1] + 1 (largest duration of the chunk)
        uint32_t* h_seg_counts = nullptr;   // pinned
        uint32_t seg_max = 0;
        unsigned int nt = 0;                // ticks inside the chunk
        bool external = false;              // the marks came from the host (apm_set_tick_marks), not from the tick finder
        int64_t latest0 = 0;                // latestBucket before the chunk
        cudaEvent_t ev = nullptr;           // the tick marks are on the host
    };
    PrePass pre[2];
    cudaStream_t s_pre = nullptr;
    uint64_t tf_tiles_cap = 0;
    uint32_t max_ticks_per_chunk = 4096;
    // ticks the host has already located (apm_set_tick_marks; key-sharded deployments, where a tick is a property of the
    // global stream): they replace the tick finder for the next process_* call, which consumes them chunk by chunk
    std::vector<TickMark> ext_marks;
    bool ext_armed = false;
    size_t ext_next = 0;        // first mark not yet handed to a chunk
    uint64_t ext_off = 0;       // offset, inside the call, of the chunk whose pre-pass is being started
    bool ext_last = false;      // ... and whether it is the call's last chunk (takes the marks at pos == n)
    uint64_t host_count[APM_MAX_SLOTS];   // log entries per slot (== SlotMeta.count)
    uint64_t host_sorted[APM_MAX_SLOTS];  // entries that were in place at the slot's last sort
    uint32_t known_max_value = 0;         // largest duration appended to any log so far

    // --- staging of host batches: pinned SoA ring (acquire/commit) + two device buffers
    apm_tx_columns pin = {nullptr, nullptr, nullptr, nullptr};
    apm_tx_columns dev[2] = {{nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr}};
    uint64_t acquired = 0;
    uint32_t* pk[2] = {nullptr, nullptr};  // staging of 12-byte / 8-byte transactions: [3][max_batch] u32 each (apm_process_host_packed, _packed8)
    apm_tx_escape* esc[2] = {nullptr, nullptr};  // a chunk's slice of the 8-byte form's escape list
    uint32_t esc_cap[2] = {0, 0};

    // --- tick outputs
    apm_stat_record* d_stats = nullptr;  // [out_ticks][K]
    apm_fs_record* d_fs = nullptr;       // [out_ticks][num_lags][K]
    apm_fs_record* d_alerts = nullptr;   // [out_ticks][max_alerts_per_tick]
    unsigned int* d_alert_count = nullptr;  // [out_ticks]
    uint32_t* d_alert_blocks = nullptr;     // [ceil(K * num_lags / 256)] hits per CTA of the ordered compaction
    uint32_t* k4_done = nullptr;            // inside d_alert_blocks: CTAs of baseline_few_kernel that have written their records
    double* zin[2] = {nullptr, nullptr};     // double-buffered [4][K]: tpm', val0, val1, val2 (quantised)
    uint8_t* emit[2] = {nullptr, nullptr};   // [K] key emitted a StatEntry this tick (snapshot of `seen`)
    // optional: fs records stream to a caller-owned pinned buffer as each tick's baseline finishes
    apm_fs_record* h_fs_bound = nullptr;
    uint64_t h_fs_bound_cap = 0;
    cudaStream_t s_d2h = nullptr;
    cudaEvent_t ev_d2h = nullptr;
    uint32_t ticks_buffered = 0;
    uint32_t alerts_gather_from = 0;     // first buffered tick not yet all-gathered

    // --- rolling baseline
    double* ring[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};        // [L][ring_pitch], column = series*K + key
    uint32_t* ring_len[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};  // [3][K]
    uint32_t* ring_next[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr}; // [3][K] slot the column's next push belongs in
    size_t ring_pitch = 0;  // doubles per ring row: 3K rounded up to 32 (TMA box width, 256-byte rows)
    CUtensorMap ring_tmap[APM_MAX_LAGS];  // 2-D view [L][ring_pitch] of each ring for the bulk-tensor loads
    CUtensorMap ring_tmap_few[APM_MAX_LAGS];  // the same view with a 96-slot box (baseline_few_kernel)
    bool ring_tmap_ok = false;
    uint64_t misalign_mark = 0;  // tick_seq at the last list load: the pending count is only meaningful after a later tick
    bool may_misalign = false;  // lists were loaded for keys that may be silent: run the realign kernel until none is left
    uint32_t ring_head[APM_MAX_LAGS] = {0, 0, 0, 0};                          // slot the next push writes
    double* thr[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};         // [K]
    double* infl[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};        // [K]
    // transaction passthrough (APM_F_PASSTHROUGH): pending words, due words, sort spare, per-tick output
    uint64_t* pt_pend[2] = {nullptr, nullptr};
    uint64_t* pt_emit = nullptr;
    uint64_t* pt_tmp = nullptr;
    uint64_t* pt_out = nullptr;
    unsigned long long* pt_cursors = nullptr;  // [0] pending, [1] due, [2] kept
    uint64_t pt_cap = 0, pt_out_cap = 0, pt_out_used = 0, pt_seq = 0;
    int pt_cur = 0;
    int64_t pt_base_ms = INT64_MIN;
    std::vector<uint32_t> pt_tick_counts;
    // guarded incremental baseline (APM_F_INCREMENTAL): running statistics per column
    double* inc_sum[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};
    double* inc_comp[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t* inc_cnt[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t* inc_nz[APM_MAX_LAGS] = {nullptr, nullptr, nullptr, nullptr};
    bool inc_dirty[APM_MAX_LAGS] = {false, false, false, false};  // lists were loaded: rebuild before the next tick
    // alert counter core (stream_process_alerts.js:348-434)
    apm_alert_rules alert_rules = {};
    uint32_t* alert_count = nullptr;   // [num_lags][K] recentAlertCounts
    double* alert_hard_max = nullptr;  // [K] per-key hardMaxMsAlertThreshold, <= 0: default
    uint8_t* alert_suppressed = nullptr;  // [K]

    // stats-only whole-chunk path (k1_chunk.cu): per-segment accumulator planes, first-seen segments, tick descriptors
    bool stats_chunk = true;               // APM_STATS_CHUNK=0: one reduce / clear / window launch per tick even in stats-only mode
    uint64_t stats_chunk_min_n = 262144;   // below this many transactions, or this many per segment, the per-tick path is as good
    uint64_t stats_chunk_min_seg = 8192;   // (APM_STATS_CHUNK_MIN_N / APM_STATS_CHUNK_MIN_SEG: tests lower them)
    bool stats_chunk_direct = true;  // APM_STATS_CHUNK_DIRECT=0: the hashed table also when every key fits shared memory
    void* chunk_planes = nullptr;
    size_t chunk_cells = 0;
    uint32_t* chunk_first_seg = nullptr;
    void* chunk_desc = nullptr;

    DevCounters* d_counters = nullptr;
    DevCounters* h_counters = nullptr;  // pinned mirror
    apm_counters counters;
    uint64_t launches = 0;
    uint32_t k1_tiles_per_cta = 3;  // tuning knob (APM_K1_TPC)
    int k4_stages = 0;              // APM_K4_STAGES: bulk-tensor loads in flight per CTA (2, 3, 4, 6, or 8 = the three-box instance for few lists); 0 = by grid size
    bool k4_force_ldg = false;      // APM_K4_LDG=1: register-pipelined loads instead of TMA (A/B measurements)
    bool k4_few = true;             // APM_K4_FEW=0: per-lag launches of baseline_tma_kernel also when all lags' column groups fit the SMs
    int k4_few_waves = 2;           // APM_K4_FEW_WAVES: the few-list walk is used while its CTAs fit this many waves of one CTA per SM

    // log feeder (k6_logparse.cu), created by the first apm_dict_set / apm_parse_log_host
    void* lp = nullptr;
    // multi-GPU
    void* nccl_comm = nullptr;
    int rank = 0, nranks = 1;
    // alert all-gather (comm.cu): one fixed-size block {header, records[gather_cap]} per rank and collective
    uint32_t gather_cap = 0;
    apm_fs_record* d_gather_send = nullptr;   // [1 + gather_cap]
    apm_fs_record* d_gather_recv = nullptr;   // [nranks][1 + gather_cap]
    apm_fs_record* d_gather_stage = nullptr;  // [nranks * gather_cap]: unpacked records for callers without a pinned buffer
    void* d_gather_state = nullptr;           // GatherState: records appended so far, error bits
    void* h_gather_state = nullptr;           // pinned copy, refreshed behind every collective
    cudaStream_t s_comm = nullptr;            // pack -> ncclAllGather -> unpack, beside the baseline stream
    cudaEvent_t ev_comm = nullptr;
    apm_fs_record* h_gather_bound = nullptr;  // apm_bind_gather_output: every tick's alerts are gathered as the tick finishes
    uint64_t h_gather_bound_cap = 0;

    // optional per-kernel-group timing (CUDA events on s_main around each launch)
    bool prof_on = false;
    struct ProfSpan { int group; cudaEvent_t a, b; };
    std::vector<ProfSpan> prof_spans;
    std::vector<cudaEvent_t> prof_pool;
    cudaEvent_t timer[2] = {nullptr, nullptr};

    bool device_error = false;
    std::string err;
};

enum { APM_PG_TICKFIND = 0, APM_PG_REDUCE, APM_PG_CLEAR, APM_PG_SORT, APM_PG_TICKSTATS, APM_PG_BASELINE, APM_PG_ALERTS, APM_PG_REALIGN, APM_PG_PASSTHROUGH, APM_PG_COUNT };

// Brackets the launches issued during its lifetime with an event pair when profiling is on.
struct ProfScope {
    apm_ctx* c; int idx = -1; cudaStream_t st;
    ProfScope(apm_ctx* ctx, int group, cudaStream_t stream = nullptr) : c(ctx), st(stream ? stream : ctx->s_main) {
        if (!c->prof_on) return;
        cudaEvent_t ev[2];
        for (int i = 0; i < 2; ++i) {
            if (!c->prof_pool.empty()) { ev[i] = c->prof_pool.back(); c->prof_pool.pop_back(); }
            else if (cudaEventCreate(&ev[i]) != cudaSuccess) return;
        }
        cudaEventRecord(ev[0], st);
        c->prof_spans.push_back({group, ev[0], ev[1]});
        idx = (int)c->prof_spans.size() - 1;
    }
    ~ProfScope() { if (idx >= 0) cudaEventRecord(c->prof_spans[idx].b, st); }
};

// k1_ingest.cu
apm_status apm_prepass_begin(apm_ctx* c, apm_ctx::PrePass& pp, const apm_tx_columns& d, uint64_t n, int64_t latest0, cudaStream_t st);
apm_status apm_prepass_finish(apm_ctx* c, apm_ctx::PrePass& pp, const apm_tx_columns& d, uint64_t n, cudaStream_t st);
apm_status apm_prepass_segment(apm_ctx* c, apm_ctx::PrePass& pp, const apm_tx_columns& d, uint64_t n, int64_t latest0, cudaStream_t st);
apm_status apm_launch_reduce(apm_ctx* c, const apm_tx_columns& d, uint64_t begin, uint64_t end);
// k1_chunk.cu
bool apm_stats_chunk_eligible(const apm_ctx* c, uint64_t n, uint32_t nt);
apm_status apm_stats_chunk(apm_ctx* c, const apm_tx_columns& d, uint64_t n, const apm_ctx::PrePass& pp, apm_tick_result* res);
// k3_sort.cu
apm_status apm_sort_slot_n(apm_ctx* c, int slot, uint64_t n, uint32_t max_value);
uint64_t* apm_radix_sort_words(cudaStream_t st, uint64_t* a, uint64_t* b, uint64_t n, uint32_t* hist, uint32_t* totals,
                               int vbits, int kbits, uint64_t* launches);
uint64_t apm_radix_sort_tiles(uint64_t n);
size_t apm_onesweep_scratch_bytes(uint64_t max_n);
uint64_t* apm_onesweep_sort(cudaStream_t st, uint64_t* a, uint64_t* b, uint64_t n, unsigned char* scratch, size_t scratch_bytes,
                            int vbits, int kbits, uint64_t* launches, cudaError_t* err);
apm_status apm_acc_from_sorted(apm_ctx* c, int slot, uint64_t n);
bool apm_small_tick_ok(const apm_ctx* c, uint64_t largest);
apm_status apm_small_tick(apm_ctx* c, const int* slots, const uint64_t* n, int nslots, uint64_t clear_mask);
int apm_bits_of(uint64_t v);
// k3_groupsort.cu
size_t apm_groupsort_scratch_bytes(uint32_t K);
bool apm_groupsort(cudaStream_t st, apm_groupsort_state* gs, unsigned char* scratch, uint32_t K, uint64_t* src, uint64_t* dst, uint64_t n,
                   int vbits, uint64_t* launches, cudaError_t* err);
cudaError_t apm_groupsort_plan_from(cudaStream_t st, apm_groupsort_state* gs, unsigned char* scratch, uint32_t K, const uint64_t* sorted,
                                    uint64_t n, uint64_t* launches);
// k2_tick.cu
apm_status apm_launch_clear_slots(apm_ctx* c, uint64_t slot_mask);
apm_status apm_launch_tick_stats(apm_ctx* c, const WindowSlots& w, int64_t tick_ts, uint32_t tick_slot, int buf);
// k4_baseline.cu
apm_status apm_launch_baseline(apm_ctx* c, int lag_idx, int64_t tick_ts, uint32_t tick_slot, int buf, cudaStream_t st);
bool apm_baseline_few_ok(const apm_ctx* c);
apm_status apm_launch_baseline_few(apm_ctx* c, int64_t tick_ts, uint32_t tick_slot, int buf, int alert_slot, cudaStream_t st);
apm_status apm_baseline_make_tmaps(apm_ctx* c);
apm_status apm_launch_realign(apm_ctx* c, int buf, cudaStream_t st);
apm_status apm_launch_alert_counter(apm_ctx* c, uint32_t fs_slot, cudaStream_t st);
apm_status apm_launch_compact_alerts(apm_ctx* c, uint32_t fs_slot, uint32_t alert_slot, cudaStream_t st);
// k5_passthrough.cu
apm_status apm_pt_append(apm_ctx* c, const apm_tx_columns& d, uint64_t begin, uint64_t end);
apm_status apm_pt_tick(apm_ctx* c, int64_t edge_ts, uint32_t tick_slot);
// k6_logparse.cu
void apm_lp_destroy(apm_ctx* c);
// comm.cu
void apm_comm_destroy_internal(apm_ctx* c);
apm_status apm_comm_tick_gather(apm_ctx* c, uint32_t alert_slot, cudaStream_t after);  // bound mode: this tick's alerts
apm_status apm_comm_release(apm_ctx* c);
// capi.cu
apm_status apm_flush_pending(apm_ctx* c);
