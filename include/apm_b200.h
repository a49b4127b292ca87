/*
 * apm_b200.h - C ABI of libapm_b200.so: the B200-native replacement for the
 * arithmetic of APMBackend's two hot stream stages,
 *     stream_calc_stats.js   (per-interval transaction-timing aggregation)
 *     stream_calc_z_score.js (rolling baseline / signal)
 * All "replaces" citations are file:line in the reference tree.
 *
 * Conventions
 *   - plain C types only; every call returns apm_status (0 = OK) and never
 *     throws or aborts.  apm_last_error(ctx) gives the text of the last failure.
 *   - one caller thread per context (the reference stages are single-threaded
 *     event loops, stream_calc_stats.js:331-371).
 *   - the context owns all device and pinned memory; output buffers passed to
 *     apm_fetch_* are caller-owned host memory.
 *   - there is no CPU fallback: without a CUDA device every compute entry
 *     point fails with APM_E_DEVICE.
 */
#ifndef APM_B200_H
#define APM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define APM_ABI_VERSION 1

typedef int32_t apm_status;
enum {
    APM_OK = 0,
    APM_E_INVALID = 1,    /* bad argument / bad config                                  */
    APM_E_DEVICE = 2,     /* CUDA error (sticky: destroy the context, state_load again) */
    APM_E_CAPACITY = 3,   /* a configured capacity was exceeded (see apm_last_error)     */
    APM_E_WOULDBLOCK = 4, /* output ring full: fetch results, then retry                */
    APM_E_STATE = 5,      /* call order violated (e.g. commit without acquire)          */
    APM_E_IO = 6,         /* state file could not be read / written                     */
    APM_E_COMM = 7        /* NCCL failure                                               */
};

#define APM_MAX_PCT 4
#define APM_MAX_LAGS 4
#define APM_NUM_SERIES 3 /* average + two percentiles, as FullStatEntry has (entries.js:86-108) */

/* apm_config.flags */
#define APM_F_PERCENTILES 1u /* keep per-bucket value logs and compute percentiles */
#define APM_F_ZSCORE 2u      /* run the rolling-baseline stage                      */
#define APM_F_KEEP_STATS 4u  /* retain apm_stat_record per tick for apm_fetch_stats */
#define APM_F_KEEP_FS 8u     /* retain apm_fs_record per tick for apm_fetch_fs      */
#define APM_F_ALERTS 16u     /* compact signalled fs records for apm_fetch_alerts / all-gather */
#define APM_F_PASSTHROUGH 64u /* keep a device-side reorder buffer and emit the time-ordered transaction passthrough at
                                 every tick (replaces the stats stage's min-heap) */
#define APM_F_INCREMENTAL 32u /* guarded incremental baseline: running sums instead of the exact-order rescan of every
                                 list; signals and lists stay bit-identical (exact re-sum inside a 1e-9 guard band),
                                 rolling mean / bounds agree to ~1e-12 relative instead of bit for bit */

/* apm_config.sd_mode */
#define APM_SD_REFERENCE 0 /* "stddev" = sqrt(mean), as util_methods.js:26-50 computes it */
#define APM_SD_TRUE 1      /* extension: population stddev of the valid ring entries       */

/*
 * Replaces the resolved numeric settings both stages read from
 * config/apm_config.json:105-107 (interval, window, buffer) and :126-172
 * (LAG / THRESHOLD / INFLUENCE defaults).  Per-key overrides are resolved by
 * the host with the reference's own logic (stream_calc_z_score.js:106-150)
 * and pushed with apm_set_key_params.
 */
typedef struct apm_config {
    uint32_t struct_size;  /* sizeof(apm_config), for ABI checking                         */
    int32_t device;        /* CUDA device ordinal                                           */
    uint32_t num_services; /* capacity of service ids                                       */
    uint32_t num_jvms;     /* capacity of JVM (server) ids; key = jvm*num_services+service  */
    int64_t bucket_ms;     /* 10000 in the reference (generateBucketLabel, :89-91)          */
    int32_t window_buckets; /* 30: windowSizeInIntervals                                    */
    int32_t buffer_buckets; /* 6:  bufferSizeInIntervals                                    */
    int32_t num_percentiles;           /* 0..APM_MAX_PCT; the reference has {75, 95}        */
    int32_t z_series_pct[2];           /* which percentiles feed z-series 1 and 2 ({0,1})   */
    double percentiles[APM_MAX_PCT];
    int32_t num_lags;                  /* 0..APM_MAX_LAGS; the reference has {360, 8640}    */
    int32_t lags[APM_MAX_LAGS];        /* ascending                                         */
    double default_threshold[APM_MAX_LAGS];
    double default_influence[APM_MAX_LAGS];
    int32_t sd_mode;                   /* APM_SD_*                                          */
    uint32_t flags;                    /* APM_F_*                                           */
    uint64_t max_tx_per_bucket;        /* capacity of one bucket's value log (tx)           */
    uint64_t max_batch;                /* capacity of the pinned ingest ring (tx)           */
    uint32_t out_ticks;                /* ticks of retained records before WOULDBLOCK       */
    uint32_t max_alerts_per_tick;      /* capacity of the compact alert block               */
} apm_config;

/* Columnar transaction batch: the fields of TxEntry (entries.js:1-21) that feed
 * statistics - server (as jvm_id), service (as service_id), endTs, elapsed.
 * Preconditions (checked on the device, violations are counted and skipped):
 * service_id < num_services, jvm_id < num_jvms, elapsed_ms >= 0, end_ts_ms >= 0. */
typedef struct apm_tx_columns {
    uint32_t* service_id;
    uint32_t* jvm_id;
    int64_t* end_ts_ms;
    int32_t* elapsed_ms;
} apm_tx_columns;

/* One per (key, tick): what generateAllStatsToQueue (stream_calc_stats.js:157-203)
 * puts into a StatEntry, un-rounded, plus the extension fields. 96 bytes. */
typedef struct apm_stat_record {
    int64_t tick_ts;  /* edgeTs = (latest - buffer - 1) * bucket_ms (:356)             */
    uint32_t key;
    uint32_t flags;   /* bit0: key has been seen (record valid); bit1: window non-empty */
    uint64_t count;   /* windowCntSum                                                   */
    uint64_t sum;     /* windowRtTotalSum (exact)                                       */
    double tpm;       /* count / (window * interval_s / 60)                             */
    double avg;       /* sum / count, NaN when the window is empty ("undefined")        */
    double pct[APM_MAX_PCT]; /* calcPercentile results, NaN when empty / not configured */
    int32_t min;      /* extension (0 when empty)                                       */
    int32_t max;      /* extension                                                      */
    double stddev;    /* extension: population stddev of the window, NaN when empty     */
} apm_stat_record;

/* One per (key, lag, tick): the numeric content of a FullStatEntry
 * (entries.js:86-118) as processData builds it (stream_calc_z_score.js:195-311).
 * NaN stands for JS `undefined`. 128 bytes. */
typedef struct apm_fs_record {
    int64_t tick_ts;
    uint32_t key;
    uint32_t lag;
    double tpm;                  /* parseFloat(tpm.toFixed(2))                           */
    double val[APM_NUM_SERIES];  /* parseFloat(x.toFixed(1)) of average, pctA, pctB       */
    double avg[APM_NUM_SERIES];  /* rolling mean of the lag list                          */
    double lb[APM_NUM_SERIES];
    double ub[APM_NUM_SERIES];
    int8_t signal[APM_NUM_SERIES];
    uint8_t flags;               /* bit0: record valid; bit1: triggerAlert (alert rules on) */
    uint8_t pad_[4];             /* with alert rules on: [0] causes (APM_CAUSE_*), [1] bad-interval count (<= 255); gathered records: [2] sending rank */
} apm_fs_record;

/* causes of an alert, in the order processFSEntry tests them (stream_process_alerts.js:398-423) */
#define APM_CAUSE_AVG_HARD_MAX 1u  /* 'average exceeded hard ms threshold' */
#define APM_CAUSE_P75_HARD_MAX 2u  /* 'per75 exceeded hard ms threshold'   */
#define APM_CAUSE_AVG_UB 4u        /* 'average UB exceeded'                */
#define APM_CAUSE_P75_UB 8u        /* 'per75 UB exceeded'                  */
#define APM_CAUSE_BOTH_UB 16u      /* 'average and per75 UB exceeded'      */

/* The deterministic core of the alert stage: replaces the counter logic of
 * processFSEntry (stream_process_alerts.js:348-434) for every (key, lag) of a tick, fused behind
 * the baseline kernels.  The wall-clock cooldown, batching and e-mail (:436-468) stay on the host.
 * Values mirror config/apm_config.json:189-204. */
typedef struct apm_alert_rules {
    uint32_t struct_size;
    int32_t enabled;
    int32_t window;              /* rollingAlertWindowSizeInIntervals (60)                     */
    int32_t required;            /* requiredNumberBadIntervalsInAlertWindowToTrigger (45)      */
    int32_t alert_on_both_only;  /* alertOnBothOnly                                            */
    uint32_t suppressed_lag_mask; /* bit i: lag index i is in suppressedLags                   */
    double hard_max_ms;          /* hardMaxMsAlertThreshold (per-key overrides: apm_set_alert_key_params) */
    double hard_min_ms;          /* hardMinMsAlertThreshold                                    */
    double hard_min_tpm;         /* hardMinTpmAlertThreshold                                   */
} apm_alert_rules;

typedef struct apm_tick_result {
    uint64_t ticks;        /* ticks run by this call                        */
    uint64_t tx_ingested;  /* transactions reduced into buckets             */
    uint64_t tx_dropped;   /* older than the retained buckets (evicted)     */
    uint64_t tx_invalid;   /* failed the apm_tx_columns preconditions       */
    int64_t latest_bucket; /* latestBucket after the call                   */
} apm_tick_result;

typedef struct apm_counters {
    uint64_t tx_ingested, tx_dropped, tx_invalid;
    uint64_t ticks, stat_records, fs_records, alerts;
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t kernel_launches;
    uint64_t keys_seen;
    uint64_t inc_fallbacks; /* APM_F_INCREMENTAL: columns re-summed exactly because the signal test was inside the guard band */
} apm_counters;

typedef struct apm_ctx apm_ctx;

/* ---- lifecycle (replaces `new StatParser()` / `new ZScoreParser()`,
 *      stream_calc_stats.js:225, stream_calc_z_score.js:367) ---- */
apm_status apm_ctx_create(const apm_config* cfg, apm_ctx** out);
void apm_ctx_destroy(apm_ctx* ctx);
const char* apm_last_error(const apm_ctx* ctx); /* ctx may be NULL (create failures) */
uint32_t apm_abi_version(void);

/* ---- ingest: replaces addData (stream_calc_stats.js:115-134) ----
 * acquire/commit: zero-copy pinned SoA slices the host parser fills; commit
 * copies them to the device on the copy stream and reduces them into their
 * (key, bucket) accumulators.  All n transactions belong to the current
 * segment: the caller has already run apm_tick for any larger bucket label,
 * exactly as consumeMsg does (:348-370). */
apm_status apm_ingest_acquire(apm_ctx* ctx, uint64_t n, apm_tx_columns* out_pinned);
apm_status apm_ingest_commit(apm_ctx* ctx, uint64_t n);

/* ---- tick: replaces the body of `if (bucketLabel > data.latestBucket)`
 * (stream_calc_stats.js:348-367: removeOldBuckets + generateAllStatsToQueue)
 * and, fused behind it, processData for every emitted StatEntry
 * (stream_calc_z_score.js:195-311). latest_bucket must exceed the current one. */
apm_status apm_tick(apm_ctx* ctx, int64_t latest_bucket, apm_tick_result* res);

/* ---- stream chunk: replaces a run of consumeMsg calls (:331-371) over n
 * transactions in arrival order, ticks found on the device.  *_host takes
 * host columns (pageable or pinned) and includes the H2D copies; *_device
 * takes device-resident columns. res may be NULL.  A call longer than max_batch goes through in
 * chunks; when a chunk holds more ticks than the output ring has room for, the call stops there with
 * APM_E_WOULDBLOCK and res counts what the chunks before it consumed: after a fetch + apm_release_ticks
 * the caller resubmits from transaction tx_ingested + tx_dropped + tx_invalid on. */
apm_status apm_process_host(apm_ctx* ctx, const apm_tx_columns* host_cols, uint64_t n, apm_tick_result* res);
apm_status apm_process_device(apm_ctx* ctx, const apm_tx_columns* dev_cols, uint64_t n, apm_tick_result* res);

/* The same run of consumeMsg calls from 12-byte instead of 20-byte transactions, for hosts whose ingest is
 * bound by the PCIe link: key = jvm_id << 16 | service_id (both below 65536), end_ts_delta_ms = endTs -
 * base_ts_ms (the chunk spans less than 2^32 ms = 49 days), elapsed_ms as unsigned (a value above INT32_MAX is
 * the invalid transaction a negative elapsed is in the wide form).  Widened on the device, then identical to
 * apm_process_host: same ticks, same records. */
typedef struct apm_tx_packed {
    const uint32_t* key;
    const uint32_t* end_ts_delta_ms;
    const uint32_t* elapsed_ms;
} apm_tx_packed;
apm_status apm_process_host_packed(apm_ctx* ctx, const apm_tx_packed* host_cols, int64_t base_ts_ms, uint64_t n,
                                   apm_tick_result* res);

/* The same run once more from 8-byte transactions, for hosts that feed several GPUs through one PCIe root / memory
 * system: one word per transaction,
 *     bits 63..48  key = jvm_id * num_services + service_id (below 65536; a value >= the context's key count is the
 *                  invalid transaction an out-of-range id is in the wide form)
 *     bits 47..24  endTs - base_ts_ms in ms (the call spans less than 2^24 ms = 4.6 h)
 *     bits 23..0   elapsed in ms; 0xFFFFFF = "look it up": the transaction's index (inside this call) is in `escapes`,
 *                  ascending by index, which carries the full int32 (a duration of 16 777 215 ms or more, or a negative,
 *                  i.e. invalid, one).  An all-ones field without a list entry stands for 16 777 215 ms itself.
 * Widened on the device, then identical to apm_process_host: same ticks, same records, same counters. */
typedef struct apm_tx_escape {
    uint32_t index;
    int32_t elapsed_ms;
} apm_tx_escape;
apm_status apm_process_host_packed8(apm_ctx* ctx, const uint64_t* tx, const apm_tx_escape* escapes, uint32_t num_escapes,
                                    int64_t base_ts_ms, uint64_t n, apm_tick_result* res);

/* ---- ticks located by the host: replaces, for the NEXT apm_process_host / _host_packed / _host_packed8 / _device call only,
 * the `bucketLabel > data.latestBucket` test of consumeMsg (stream_calc_stats.js:348).  In a key-sharded
 * deployment a tick is a property of the GLOBAL stream - any transaction with a larger bucket label triggers
 * it, whichever rank owns its key - so the host that bins the transactions by owner (apm_partition_host)
 * tells every rank where the ticks fall in its own sub-stream; a rank then ticks even while its shard is silent.
 * marks[i]: the tick for bucket `label` fires before the transaction at index `pos` of the next call
 * (pos == n: behind its last transaction).  Ascending in pos, strictly ascending in label, above latestBucket.
 * With marks set the device does not look for ticks itself; n == 0 arms a call without any tick.  The marks are
 * consumed by that call whether it succeeds or not (after APM_E_WOULDBLOCK: set the remaining ones again,
 * rebased, before resubmitting). */
typedef struct apm_tick_mark {
    uint64_t pos;
    int64_t label;
} apm_tick_mark;
apm_status apm_set_tick_marks(apm_ctx* ctx, const apm_tick_mark* marks, uint32_t n);

/* ---- per-key z-score settings: replaces updateServiceSettings
 * (stream_calc_z_score.js:134-150); thr/infl hold n values for keys
 * [key_lo, key_lo+n) of lag index lag_idx. */
apm_status apm_set_key_params(apm_ctx* ctx, uint32_t key_lo, uint32_t n, int32_t lag_idx,
                              const double* thr, const double* infl);
/* With rules enabled every fs record carries triggerAlert / causes / the bad-interval count, and the
 * compact alert block (apm_fetch_alerts, apm_gather_alerts) holds the records with triggerAlert set
 * instead of every record with a signal.  hard_max_ms[i] <= 0 keeps the default for key key_lo + i
 * (the reference's `!= 0` test, :396); suppressed[i] != 0 puts the key's service in suppressedServices. */
apm_status apm_set_alert_rules(apm_ctx* ctx, const apm_alert_rules* rules);
apm_status apm_set_alert_key_params(apm_ctx* ctx, uint32_t key_lo, uint32_t n, const double* hard_max_ms,
                                    const uint8_t* suppressed);
/* replaces removeStaleLagData (:170-193) for one lag: forget its lists. */
apm_status apm_drop_lag(apm_ctx* ctx, int32_t lag_idx);

/* ---- results: replace the outQueue.writeLineToQueue calls (stream_calc_stats.js:200,
 * stream_calc_z_score.js:473) with binary records.  The records of the ticks
 * run since the last apm_release_ticks are retained on the device (up to
 * out_ticks ticks); a fetch copies the valid ones, tick-major, key-ascending
 * inside a tick (fs: lag-major then key; the host re-orders to first-seen
 * order, the order Object.entries gives the reference, :159-160).  cap must
 * hold every retained record slot: ticks * keys (* lags for fs).  Alerts are
 * the fs records with a non-zero signal, (lag, key)-ascending inside a tick.
 * apm_release_ticks frees the retained ticks for reuse. */
apm_status apm_fetch_stats(apm_ctx* ctx, apm_stat_record* out, uint64_t cap, uint64_t* n);
apm_status apm_fetch_fs(apm_ctx* ctx, apm_fs_record* out, uint64_t cap, uint64_t* n);
apm_status apm_fetch_alerts(apm_ctx* ctx, apm_fs_record* out, uint64_t cap, uint64_t* n);
apm_status apm_release_ticks(apm_ctx* ctx);
/* With APM_F_PASSTHROUGH: replaces writeHeapToQueue (stream_calc_stats.js:136-155).  Every valid
 * transaction that was ingested gets the next arrival sequence number (counted over all submitted
 * transactions, modulo 2^24); at each tick those with endTs <= edgeTs leave the device's reorder
 * buffer in non-decreasing endTs order (ties in arrival order).  Returns the retained ticks' output
 * concatenated (end_ts and/or seq may be NULL) and how many entries each tick released. */
apm_status apm_fetch_passthrough(apm_ctx* ctx, int64_t* end_ts, uint32_t* seq, uint64_t cap, uint64_t* n,
                                 uint32_t* tick_counts, uint32_t tick_cap, uint32_t* nticks);
/* Page-locked host memory for the `out` buffers above (the N-API addon wraps it as an external
 * ArrayBuffer): a fetch into it is one DMA instead of a staged copy.  No reference counterpart. */
apm_status apm_host_alloc(uint64_t bytes, void** out);
/* Streams the fs records of every tick into `pinned` (apm_host_alloc memory, cap_records >= out_ticks *
 * keys * lags) as soon as that tick's baseline has finished, under the following ticks' work; a later
 * apm_fetch_fs(ctx, pinned, ...) then only waits for the last copy.  NULL unbinds. */
apm_status apm_bind_fs_output(apm_ctx* ctx, apm_fs_record* pinned, uint64_t cap_records);
void apm_host_free(void* p);
uint32_t apm_ticks_buffered(const apm_ctx* ctx);
/* wait for all queued device work (for timing) */
apm_status apm_sync(apm_ctx* ctx);

/* ---- lag lists: replaces the avgList/per75List/per95List arrays a resume
 * file carries (stream_calc_z_score.js:37-64).  values holds `len` entries,
 * oldest first, NaN for null; len <= lag. */
apm_status apm_ring_load(apm_ctx* ctx, uint32_t key, int32_t lag_idx, int32_t series,
                         const double* values, uint32_t len);
apm_status apm_ring_read(apm_ctx* ctx, uint32_t key, int32_t lag_idx, int32_t series,
                         double* values, uint32_t cap, uint32_t* len);
/* all keys at once: values[len][APM_NUM_SERIES][keys], oldest row first, host or device memory. */
apm_status apm_ring_load_all(apm_ctx* ctx, int32_t lag_idx, const double* values, uint32_t len, int32_t on_device);

/* ---- checkpoint: replaces saveToResumeFile / resumeFromFileIfExists
 * (stream_calc_stats.js:54-87, stream_calc_z_score.js:37-64) with a compact
 * binary image of the same state: bucket accumulators and value logs, seen keys, lag lists and per-key
 * settings, alert counters and - with APM_F_PASSTHROUGH - the transactions still waiting in the device-side
 * reorder buffer with their sequence numbers (the `minHeap` the reference's stats resume file carries,
 * :40-52); without that flag the heap is the host's to save.  Load needs a context created with the same
 * configuration; retained output records are not part of the image (fetch them before saving). */
apm_status apm_state_save(apm_ctx* ctx, const char* path);
apm_status apm_state_load(apm_ctx* ctx, const char* path);

/* ---- optional device timing (no reference counterpart; the reference only times its DB
 * inserts, stream_insert_db.js:295-308).  When enabled, every launch group is bracketed by
 * CUDA events on the context's stream; read sums them per group and resets.
 * groups: 0 chunk pre-pass (tick finder + segment table), 1 keyed reduce, 2 slot clear, 3 radix sort, 4 window/percentiles,
 *         5 rolling baseline, 6 alert compaction, 7 lag-list realignment, 8 transaction passthrough. */
#define APM_KT_GROUPS 12
typedef struct apm_kernel_times {
    double ms[APM_KT_GROUPS];
    uint64_t spans[APM_KT_GROUPS]; /* event-bracketed launch groups (one per tick and kernel group) */
} apm_kernel_times;
apm_status apm_profile_enable(apm_ctx* ctx, int32_t on);
/* The baseline of tick i normally runs on a second stream under the ingest / sort of tick i+1;
 * 0 runs everything in line on one stream (isolated kernel timings).  Results are identical. */
apm_status apm_set_overlap(apm_ctx* ctx, int32_t on);
/* stopwatch: CUDA events on the context's stream; stop waits for the queued work and returns ms. */
apm_status apm_timer_start(apm_ctx* ctx);
apm_status apm_timer_stop(apm_ctx* ctx, double* ms);
apm_status apm_profile_read(apm_ctx* ctx, apm_kernel_times* out);

/* ---- interop with the reference's JSON resume files (the host converts the schemas of
 * stream_calc_stats.js:54-87 and stream_calc_z_score.js:37-64, apmbackend_b200/resume_interop.py) ----
 * apm_set_latest: latestBucket of a resumed StatParser without running a tick (retained bucket slots are
 *   assigned to the labels latest-keep .. latest and emptied).
 * apm_mark_seen: keys that exist in `servers{}` even if none of their buckets is retained.
 * apm_log_read: the raw (key << 32 | elapsed) words of one retained bucket, i.e. buckets[label] of every key. */
apm_status apm_set_latest(apm_ctx* ctx, int64_t latest_bucket);
apm_status apm_mark_seen(apm_ctx* ctx, uint32_t key_lo, uint32_t n);
apm_status apm_log_read(apm_ctx* ctx, int64_t label, uint64_t* out, uint64_t cap, uint64_t* n);
int64_t apm_latest(const apm_ctx* ctx);

/* Loads every kernel of the library (and the static scratch of apm_sorted_percentiles_noalloc) on `device` now
 * instead of at first launch.  CUDA's default lazy loading makes the first launch of each kernel a module load that
 * may need device memory and a drained device: a host that reserves all GPU memory for its own pool calls this
 * first, and a multi-GPU host needs it before collectives are in flight (a load cannot complete while an NCCL
 * kernel spins on a peer that is itself waiting for a load) - apm_comm_init calls it.  No reference counterpart. */
apm_status apm_preload_kernels(int32_t device);

apm_status apm_get_counters(apm_ctx* ctx, apm_counters* out);
apm_status apm_reset_counters(apm_ctx* ctx);

/* ---- exact helpers the host formatter needs (replace Number.prototype.toFixed
 * as used by nf(), entries.js:65-69): parseFloat(x.toFixed(digits)), digits in {1,2}. */
double apm_quantize(double x, int32_t digits);

/* ---- stand-alone operator: replaces Array.prototype.binaryConcat + calcPercentile
 * (util_methods.js:102-106, 112-142) for one array of n durations (>= 0) in device memory
 * on `device`: sorts them ascending into d_sorted (device, n entries, may be NULL) and writes
 * the npct requested percentiles (NaN when n == 0) to out_pct (host).  Scratch memory is
 * allocated and freed inside the call.  Errors are reported through apm_last_error(NULL). */
apm_status apm_sorted_percentiles(int32_t device, const int32_t* d_values, uint32_t n, const double* pcts, int32_t npct,
                                  double* out_pct, int32_t* d_sorted);

/* The operator for an array in HOST memory (what a JavaScript caller of Array.prototype.calcPercentile has): copied
 * to the device, sorted there, h_sorted (may be NULL) and out_pct copied back. */
apm_status apm_sorted_percentiles_host(int32_t device, const int32_t* h_values, uint32_t n, const double* pcts, int32_t npct,
                                       double* out_pct, int32_t* h_sorted);

/* ---- the same operator for callers that must not allocate.  It has no counterpart in the
 * reference.  Some hosts of this library cannot call cudaMalloc at the point where they need
 * a percentile: a stage whose process hands all of the GPU's memory to its own pool at
 * start-up (a later cudaMalloc would simply fail), a message handler on a latency budget that
 * cannot absorb the device-wide synchronisation cudaMalloc and cudaFree imply, or a
 * deployment whose policy forbids device allocations after initialisation.  For them the scratch memory
 * (two word buffers, histograms, a few scalars; about 4.3 MB) lives in a fixed array inside
 * the library image.  A GPU holds those bytes from the moment the library's code is loaded onto it until the
 * process exits: under CUDA's default lazy loading that moment is the first call on the device (or
 * apm_preload_kernels(device)), with CUDA_MODULE_LOADING=EAGER it is the creation of the CUDA context.
 *   - n may be at most APM_NOALLOC_MAX_VALUES; larger arrays are refused with APM_E_CAPACITY
 *     (use apm_sorted_percentiles).
 *   - the call itself performs no cudaMalloc / cudaFree and calls no allocating library
 *     routine; inputs (d_values) and outputs (d_sorted on the device, out_pct on the host)
 *     are the caller's.
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress) and, under
 *     lazy loading, is what loads the kernels and the array: it needs a little free device memory and must not
 *     happen inside a stream capture.  A host that hands all GPU memory to its own pool calls
 *     apm_preload_kernels(device) - or makes one call with a small n >= 2 - BEFORE it reserves the memory;
 *     from then on calls leave cudaMemGetInfo's free figure untouched (tests/test_zz_no_alloc.py checks that).
 *   - all calls share the one array, which keeps the previous call's contents (the call
 *     clears the single word it accumulates into).  Two calls that overlap on the same GPU
 *     - from two host threads or processes sharing a context - race on it: serialise them.
 *   - both this and apm_sorted_percentiles run on the legacy default stream, which also orders them against
 *     every blocking stream of the process. */
#define APM_NOALLOC_MAX_VALUES 262144u
apm_status apm_sorted_percentiles_noalloc(int32_t device, const int32_t* d_values, uint32_t n, const double* pcts,
                                          int32_t npct, double* out_pct, int32_t* d_sorted);

/* ---- multi-GPU: keys are sharded by the host (hash or placement table); each rank owns its keys' buckets and
 * baselines outright.  The only exchange is an all-gather of the compact alert records (per tick with
 * apm_bind_gather_output, or on demand). */
uint32_t apm_owner_of_key(uint64_t global_key, uint32_t nranks);
/* Placement for skewed traffic (BASELINE config 5; no counterpart in the reference, whose stages are single processes):
 * owner_out[key] for the dense key ids 0 .. num_keys-1 from the transactions each key had in an earlier interval.
 * load(key) = count + key_cost (what holding a key costs per interval, in transactions: the exact-order baseline walks
 * every key's lists each tick whatever its traffic; 0 balances transactions only).  Heavy keys go longest-first to the
 * least loaded rank, light keys keep apm_owner_of_key unless that rank is above the mean; a key is never split and rings
 * do not migrate - choose the table when a deployment starts.  Pure host arithmetic, no device needed. */
apm_status apm_owner_table_build(const uint64_t* key_counts, uint32_t num_keys, uint32_t nranks, uint64_t key_cost,
                                 uint32_t* owner_out);
/* The other placement: balance the transactions, keep the KEY COUNT of every rank within (1 + key_slack) of the mean.
 * A rank's ring memory and its exact-order baseline walk follow its keys, its ingest / sort / PCIe time its
 * transactions; with Zipf traffic over many keys (80 k keys at 8 ranks: the top key alone is two thirds of a rank's
 * fair share) the hash leaves one rank with 1.5 x the mean.  Longest-first onto the rank with the fewest transactions
 * that still has room for a key.  Pure host arithmetic. */
apm_status apm_owner_table_build_capped(const uint64_t* key_counts, uint32_t num_keys, uint32_t nranks, double key_slack,
                                        uint32_t* owner_out);
/* Host binning for a key-sharded deployment (no counterpart in the reference, whose stats stage is one process that
 * sees every transaction; it keeps what makes sharding legal - no statistic has a cross-key term,
 * stream_calc_stats.js:159-202, stream_calc_z_score.js:197-308 - and what makes it subtle: ticks are triggered by ANY
 * transaction, :348).  Splits n arrival-ordered transactions into one sub-stream per rank (arrival order kept) and
 * gives every rank the global ticks as marks for apm_set_tick_marks.  owner_table[key] (apm_owner_table_build) or NULL
 * for apm_owner_of_key; local_id[key] (optional) is the dense id the key has on its owner - the sub-stream then carries
 * service_id = local_id, jvm_id = 0 for a context sized to the rank's own keys.  Transactions that fail the
 * apm_tx_columns preconditions go to rank 0 unchanged (it counts them as invalid); like on the device their
 * timestamp still ticks.  latest_bucket: latestBucket before these transactions; *latest_out: after.
 * Pure host arithmetic, no device needed. */
typedef struct apm_partition_out {
    apm_tx_columns cols;   /* caller-owned host columns, capacity `cap` transactions */
    uint64_t cap;
    uint64_t n;            /* out: transactions routed to this rank */
    apm_tick_mark* marks;  /* caller-owned, capacity marks_cap */
    uint32_t marks_cap;
    uint32_t num_marks;    /* out */
} apm_partition_out;
apm_status apm_partition_host(const apm_tx_columns* cols, uint64_t n, int64_t bucket_ms, int64_t latest_bucket,
                              uint32_t num_services, uint32_t num_jvms, uint32_t nranks, const uint32_t* owner_table,
                              const uint32_t* local_id, apm_partition_out* out /* [nranks] */, int64_t* latest_out);
apm_status apm_comm_unique_id(void* out128);                 /* ncclGetUniqueId */
apm_status apm_comm_init(apm_ctx* ctx, int32_t rank, int32_t nranks, const void* unique_id128);
/* Collective: every rank calls it at the same point of its stream of calls (ranks tick in step: ticks are global).
 * Gathers, from all ranks, the alerts of the retained ticks that have not been gathered yet.  Each rank contributes
 * one fixed-size block {count, records[capacity]}: the alert blocks of its ticks packed on the device, ONE
 * ncclAllGather, and a kernel that appends the valid prefix of every rank's block to `out` - page-locked host memory
 * (apm_host_alloc) is written directly, pageable memory through a device staging area.  The call waits once, at its
 * end.  Order: one block per rank (rank-ascending), inside it (tick, lag, key)-ascending.  Capacity: records per
 * block, default max_alerts_per_tick (one tick always fits); a host that lets many ticks pile up raises it with
 * apm_comm_set_capacity (itself a collective: the ranks agree on the largest value any of them asked for, as they do
 * on the default in apm_comm_init - blocks must have one size).  A block or tick that overflows is reported by every rank alike (APM_E_CAPACITY): the
 * headers travel with the blocks, so no rank is left waiting in a collective the others skipped.  Every gathered
 * record carries its sending rank in pad_[2] (ranks below 256): with dense per-rank key ids the key alone does not
 * say whose key it is. */
apm_status apm_comm_set_capacity(apm_ctx* ctx, uint32_t records_per_block);
/* Streams the gather: with a bound buffer (page-locked, cap_records) every tick's alerts are all-gathered on their own
 * stream as soon as that tick's baseline has finished, under the following ticks' work, and appended tick by tick
 * (tick-major, then rank, lag, key); apm_gather_alerts(ctx, pinned, ...) then only waits for the last one and returns
 * the count; the buffer is appended to from 0 again afterwards.  NULL unbinds.  All ranks bind, or none. */
apm_status apm_bind_gather_output(apm_ctx* ctx, apm_fs_record* pinned, uint64_t cap_records);
apm_status apm_gather_alerts(apm_ctx* ctx, apm_fs_record* out, uint64_t cap, uint64_t* n);

/* ------------------------------------------------------------------------------------------------
 * Log feeder: CommonTiming lines of a WildFly log parsed on the device.
 *
 * Replaces, for a buffer of raw log text of ONE file (one server), what stream_parse_transactions.js
 * does line by line: readLine's dispatch on the four CommonTiming patterns (:733-777),
 * parseEjbCommonTimingEntry / Exit (:378-446), parseCommonTimingEntry / Exit (:451-565: token
 * positions, the INFO-based split), convertStringDateToMs (:242-256, the standard-log form) and
 * outputRecord's `Provider[x]` -> `Provider:x` rewrite, start = end - elapsed fallback and `^S:`
 * top-level test (:264-293).  Start and Stop lines of the same (logId, service) are paired inside the
 * buffer (the reference's recordCache, :393-400 / :471-480, without its 120-s wall-clock TTL).
 *
 * Stays with the host (it reads the records below): SOAP logs and parseAppLine (account numbers),
 * the need-number cache and its TTL, Start lines whose Stop is in a later buffer (flag
 * APM_LR_OPEN_START), service names not in the dictionary yet (service_id == APM_NO_SERVICE).
 * Only ASCII white space separates tokens (JavaScript's \s also matches Unicode spaces).
 * ---------------------------------------------------------------------------------------------- */
#define APM_NO_SERVICE 0xFFFFFFFFu

enum {
    APM_LK_CT_START = 1,   /* INFO *CommonTiming::Start           -> parseCommonTimingEntry     */
    APM_LK_CT_STOP = 2,    /* INFO *CommonTiming::Stop            -> parseCommonTimingExit      */
    APM_LK_EJB_ENTRY = 3,  /* INFO *[CommonTiming] The EJB        -> parseEjbCommonTimingEntry  */
    APM_LK_EJB_EXIT = 4    /* INFO *[CommonTiming] Total time     -> parseEjbCommonTimingExit   */
};
enum {
    APM_LR_LOGID_EMPTY = 1,   /* `[]`: entries are discarded, exits become incomplete records (:386, :411, :526) */
    APM_LR_BAD_TIME = 2,      /* the date / time tokens are not seven decimal fields: ts_ms is unset           */
    APM_LR_BAD_ELAPSED = 4,   /* parseInt(elapsed) is NaN                                                      */
    APM_LR_PAIRED = 8,        /* exit: its Start was found in this buffer, start_ms is that line's time        */
    APM_LR_OPEN_START = 16,   /* entry: no Stop consumed it inside this buffer                                 */
    APM_LR_BAF_INFO = 32,     /* the line matches /\[[^ ]+] +INFO / (attemptReadAccountNumberFromBAFInfo)      */
    APM_LR_TOP_LEVEL = 64,    /* the rewritten service starts with "S:"                                        */
    APM_LR_FED = 128          /* exit: well formed and known, part of the columns apm_feed_parsed ingests      */
};

typedef struct apm_log_record {
    uint64_t line_offset;     /* of the line in the text buffer */
    uint32_t line_length;     /* without the newline */
    uint8_t kind;             /* APM_LK_* */
    uint8_t flags;            /* APM_LR_* */
    uint16_t service_length;  /* of the raw service token (without the "S:" the EJB forms prepend) */
    uint32_t service_offset;  /* of the raw service token inside the line */
    uint32_t service_id;      /* dictionary id of the rewritten name, or APM_NO_SERVICE */
    int32_t elapsed_ms;       /* exits: parseInt(elapsed) */
    uint32_t logid_length;    /* of arr[0], brackets included */
    int64_t ts_ms;            /* entries: startTs, exits: endTs (epoch ms) */
    int64_t start_ms;         /* exits: paired Start's time, else end - elapsed (:279-281) */
    uint64_t logid_hash;      /* apm_name_hash of the logId with its brackets removed */
    uint64_t name_hash;       /* apm_name_hash of the rewritten service name (what the dictionary is keyed by) */
} apm_log_record;             /* 64 bytes */

/* offset_ms (local - UTC) applies to local wall-clock times >= local_from_ms (local time counted as if it
 * were UTC); rules ascending, the first one covers everything before the second.  What `new Date(y, m, d,
 * H, M, S, ms)` resolves through the host's zone database is passed in by the host. */
typedef struct apm_tz_rule {
    int64_t local_from_ms;
    int64_t offset_ms;
} apm_tz_rule;
#define APM_MAX_TZ_RULES 8

typedef struct apm_parse_result {
    uint64_t lines;            /* newline-separated lines seen */
    uint64_t records;          /* lines recognised as one of the four forms */
    uint64_t exits;            /* Stop / Total-time lines among them */
    uint64_t fed;              /* exits that went into the transaction columns */
    uint64_t unknown_service;  /* exits held back because their name is not in the dictionary */
    uint64_t malformed;        /* exits held back for APM_LR_BAD_TIME / APM_LR_BAD_ELAPSED */
} apm_parse_result;

/* FNV-1a 64 of a byte string: the hash the device computes for service names and log ids. */
uint64_t apm_name_hash(const char* s, uint64_t len);
/* Replaces the device dictionary: rewritten service name hash -> service id (< num_services). */
apm_status apm_dict_set(apm_ctx* ctx, const uint64_t* name_hashes, const uint32_t* service_ids, uint32_t n);
/* Parses `nbytes` of log text (host memory).  file_type: 0 = application log (CommonTiming::Start / Stop
 * only), 1 = server.log (the EJB forms take precedence, :764-776).  Records come back in line order; the
 * transaction columns of the fed exits stay on the device for apm_feed_parsed.  `out` may be NULL. */
apm_status apm_parse_log_host(apm_ctx* ctx, const char* text, uint64_t nbytes, uint32_t file_type, uint32_t jvm_id,
                              const apm_tz_rule* tz, uint32_t num_tz, apm_log_record* out, uint64_t cap, uint64_t* n,
                              apm_parse_result* result);
/* Runs the hot path (as apm_process_device) over the columns the last apm_parse_log_host left on the device. */
apm_status apm_feed_parsed(apm_ctx* ctx, apm_tick_result* result);

#ifdef __cplusplus
}
#endif
#endif /* APM_B200_H */
