/*
 * apm_b200.node - Node-API shim over libapm_b200.so.  Marshalling only: every method unpacks its
 * JavaScript arguments, calls one or two entry points of include/apm_b200.h and wraps the result.
 *
 *   const { ApmB200 } = require('./apm_b200.node');
 *   const gpu = new ApmB200({ numServices, numJvms, bucketMs, window, buffer, percentiles: [75, 95],
 *                             lags: [360, 8640], thresholds: [20, 15], influences: [0.1, 0], ... });
 *   gpu.addBatch(Uint32Array svc, Uint32Array jvm, BigInt64Array endTs, Int32Array elapsed)   // addData, :115-134
 *   gpu.tick(latestBucket) -> { ticks, stats: ArrayBuffer, fs: ArrayBuffer, alerts: ArrayBuffer }  // :348-367 + processData
 *   gpu.processColumns(svc, jvm, endTs, elapsed) -> same object (ticks found on the device)
 *   gpu.setKeyParams(key, lagIndex, threshold, influence)                                   // updateServiceSettings
 *   gpu.saveState(path) / gpu.loadState(path)                                               // resume files
 *   gpu.close()
 *
 * The JS classes StatParser / ZScoreParser of stream_calc_stats.js / stream_calc_z_score.js keep
 * their method names and call these.  Errors become JS exceptions
 * carrying apm_last_error(); the stages catch and log them as they do today
 * (stream_calc_stats.js:395-397).
 *
 * Built here compile-only (no Node in this image): the napi_* symbols stay undefined until `node`
 * loads the module; tests/fake_napi.c stands in for them in the test suite.
 */
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/apm_b200.h"
#include "node_api_min.h"

typedef struct {
    apm_ctx* ctx;
    apm_config cfg;
    uint32_t keys;
    apm_fs_record* fs_pinned;      /* bindFsOutput: page-locked, fs records stream into it tick by tick */
    uint64_t fs_pinned_cap;
    apm_fs_record* gather_pinned;  /* bindGatherOutput / gatherAlerts */
    uint64_t gather_cap;
    int32_t comm_ranks;            /* > 0 once commInit ran: every result object then carries `gathered` */
} addon_obj;

static napi_value undefined(napi_env env) {
    napi_value u;
    napi_get_undefined(env, &u);
    return u;
}

static napi_value fail(napi_env env, apm_ctx* ctx, apm_status st, const char* what) {
    char msg[768];
    const char* detail = apm_last_error(ctx);
    snprintf(msg, sizeof msg, "%s: apm status %d: %s", what, (int)st, detail ? detail : "");
    napi_throw_error(env, "APM_B200", msg);
    return NULL;
}

/* ---- reading the options object ------------------------------------------------------------ */
static bool get_number(napi_env env, napi_value obj, const char* name, double* out) {
    bool has = false;
    napi_value v;
    napi_valuetype t;
    if (napi_has_named_property(env, obj, name, &has) != napi_ok || !has) return false;
    if (napi_get_named_property(env, obj, name, &v) != napi_ok) return false;
    if (napi_typeof(env, v, &t) != napi_ok || t != napi_number) return false;
    return napi_get_value_double(env, v, out) == napi_ok;
}

static int get_number_array(napi_env env, napi_value obj, const char* name, double* out, int cap) {
    bool has = false, is_arr = false;
    napi_value arr;
    uint32_t n = 0;
    if (napi_has_named_property(env, obj, name, &has) != napi_ok || !has) return -1;
    if (napi_get_named_property(env, obj, name, &arr) != napi_ok) return -1;
    if (napi_is_array(env, arr, &is_arr) != napi_ok || !is_arr) return -1;
    napi_get_array_length(env, arr, &n);
    if ((int)n > cap) return -2;
    for (uint32_t i = 0; i < n; ++i) {
        napi_value e;
        napi_get_element(env, arr, i, &e);
        if (napi_get_value_double(env, e, &out[i]) != napi_ok) return -1;
    }
    return (int)n;
}

static void finalize_obj(napi_env env, void* data, void* hint) {
    (void)env; (void)hint;
    addon_obj* o = (addon_obj*)data;
    if (o) {
        if (o->ctx) apm_ctx_destroy(o->ctx);
        apm_host_free(o->fs_pinned);
        apm_host_free(o->gather_pinned);
        free(o);
    }
}

/* new ApmB200(options) - defaults are the reference's settings (config/apm_config.json:105-107, 126-172) */
static napi_value Construct(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1], self;
    napi_get_cb_info(env, info, &argc, argv, &self, NULL);
    if (argc < 1) { napi_throw_error(env, "APM_B200", "new ApmB200(options): options object expected"); return NULL; }
    addon_obj* o = (addon_obj*)calloc(1, sizeof *o);
    apm_config* c = &o->cfg;
    double d, arr[APM_MAX_LAGS];
    c->struct_size = sizeof *c;
    c->device = get_number(env, argv[0], "device", &d) ? (int32_t)d : 0;
    c->num_services = get_number(env, argv[0], "numServices", &d) ? (uint32_t)d : 512;
    c->num_jvms = get_number(env, argv[0], "numJvms", &d) ? (uint32_t)d : 16;
    c->bucket_ms = get_number(env, argv[0], "bucketMs", &d) ? (int64_t)d : 10000;
    c->window_buckets = get_number(env, argv[0], "window", &d) ? (int32_t)d : 30;
    c->buffer_buckets = get_number(env, argv[0], "buffer", &d) ? (int32_t)d : 6;
    int n = get_number_array(env, argv[0], "percentiles", c->percentiles, APM_MAX_PCT);
    if (n == -1) { c->percentiles[0] = 75; c->percentiles[1] = 95; n = 2; }
    if (n < 0) { free(o); napi_throw_error(env, "APM_B200", "percentiles: at most 4 numbers"); return NULL; }
    c->num_percentiles = n;
    c->z_series_pct[0] = 0; c->z_series_pct[1] = 1;
    if (get_number_array(env, argv[0], "zSeriesPct", arr, 2) == 2) { c->z_series_pct[0] = (int32_t)arr[0]; c->z_series_pct[1] = (int32_t)arr[1]; }
    n = get_number_array(env, argv[0], "lags", arr, APM_MAX_LAGS);
    if (n == -1) { arr[0] = 360; arr[1] = 8640; n = 2; }
    if (n < 0) { free(o); napi_throw_error(env, "APM_B200", "lags: at most 4 numbers"); return NULL; }
    c->num_lags = n;
    for (int i = 0; i < n; ++i) c->lags[i] = (int32_t)arr[i];
    if (get_number_array(env, argv[0], "thresholds", c->default_threshold, APM_MAX_LAGS) != n) { c->default_threshold[0] = 20; c->default_threshold[1] = 15; }
    if (get_number_array(env, argv[0], "influences", c->default_influence, APM_MAX_LAGS) != n) { c->default_influence[0] = 0.1; c->default_influence[1] = 0; }
    c->sd_mode = APM_SD_REFERENCE;
    c->flags = APM_F_PERCENTILES | APM_F_ZSCORE | APM_F_KEEP_STATS | APM_F_KEEP_FS | APM_F_ALERTS;
    if (get_number(env, argv[0], "flags", &d)) c->flags = (uint32_t)d;
    c->max_tx_per_bucket = get_number(env, argv[0], "maxTxPerBucket", &d) ? (uint64_t)d : (1u << 20);
    c->max_batch = get_number(env, argv[0], "maxBatch", &d) ? (uint64_t)d : (1u << 20);
    c->out_ticks = get_number(env, argv[0], "outTicks", &d) ? (uint32_t)d : 64;
    c->max_alerts_per_tick = get_number(env, argv[0], "maxAlertsPerTick", &d) ? (uint32_t)d : 4096;
    o->keys = c->num_services * c->num_jvms;
    apm_status st = apm_ctx_create(c, &o->ctx);
    if (st) { free(o); return fail(env, NULL, st, "apm_ctx_create"); }
    napi_wrap(env, self, o, finalize_obj, NULL, NULL);
    return self;
}

static addon_obj* self_of(napi_env env, napi_callback_info info, size_t* argc, napi_value* argv) {
    napi_value self;
    addon_obj* o = NULL;
    napi_get_cb_info(env, info, argc, argv, &self, NULL);
    if (napi_unwrap(env, self, (void**)&o) != napi_ok || !o || !o->ctx) {
        napi_throw_error(env, "APM_B200", "ApmB200 object is closed");
        return NULL;
    }
    return o;
}

/* the four typed arrays of a transaction batch */
static bool get_columns(napi_env env, napi_value* argv, apm_tx_columns* cols, size_t* n) {
    napi_typedarray_type t[4];
    size_t len[4];
    void* p[4];
    for (int i = 0; i < 4; ++i)
        if (napi_get_typedarray_info(env, argv[i], &t[i], &len[i], &p[i], NULL, NULL) != napi_ok) return false;
    if (t[0] != napi_uint32_array || t[1] != napi_uint32_array || t[2] != napi_bigint64_array || t[3] != napi_int32_array) return false;
    if (len[0] != len[1] || len[0] != len[2] || len[0] != len[3]) return false;
    cols->service_id = (uint32_t*)p[0]; cols->jvm_id = (uint32_t*)p[1]; cols->end_ts_ms = (int64_t*)p[2]; cols->elapsed_ms = (int32_t*)p[3];
    *n = len[0];
    return true;
}

/* addBatch(svc, jvm, endTs, elapsed): addData for a batch without a tick inside */
static napi_value AddBatch(napi_env env, napi_callback_info info) {
    size_t argc = 4, n = 0;
    napi_value argv[4];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    apm_tx_columns src, pin;
    if (argc < 4 || !get_columns(env, argv, &src, &n)) {
        napi_throw_error(env, "APM_B200", "addBatch(Uint32Array, Uint32Array, BigInt64Array, Int32Array) of equal length");
        return NULL;
    }
    size_t off = 0;
    while (off < n) {
        size_t m = n - off < o->cfg.max_batch ? n - off : (size_t)o->cfg.max_batch;
        apm_status st = apm_ingest_acquire(o->ctx, m, &pin);
        if (st) return fail(env, o->ctx, st, "apm_ingest_acquire");
        memcpy(pin.service_id, src.service_id + off, 4 * m);
        memcpy(pin.jvm_id, src.jvm_id + off, 4 * m);
        memcpy(pin.end_ts_ms, src.end_ts_ms + off, 8 * m);
        memcpy(pin.elapsed_ms, src.elapsed_ms + off, 4 * m);
        st = apm_ingest_commit(o->ctx, m);
        if (st) return fail(env, o->ctx, st, "apm_ingest_commit");
        off += m;
    }
    return undefined(env);
}

static apm_status gather_buffer(addon_obj* o, uint64_t cap);

/* { ticks, latestBucket, stats, fs, alerts[, gathered][, passthrough] } with the records of the retained ticks, then releases them */
static napi_value collect(napi_env env, addon_obj* o, const apm_tick_result* res) {
    napi_value out, v;
    void* p;
    uint64_t n = 0;
    const uint32_t ticks = apm_ticks_buffered(o->ctx);
    napi_create_object(env, &out);
    napi_create_double(env, (double)res->ticks, &v);
    napi_set_named_property(env, out, "ticks", v);
    napi_create_int64(env, res->latest_bucket, &v);
    napi_set_named_property(env, out, "latestBucket", v);
    if (o->cfg.flags & APM_F_KEEP_STATS) {
        const uint64_t cap = (uint64_t)ticks * o->keys;
        apm_stat_record* tmp = (apm_stat_record*)malloc((cap ? cap : 1) * sizeof *tmp);
        apm_status st = apm_fetch_stats(o->ctx, tmp, cap, &n);
        if (st) { free(tmp); return fail(env, o->ctx, st, "apm_fetch_stats"); }
        napi_create_arraybuffer(env, n * sizeof *tmp, &p, &v);
        memcpy(p, tmp, n * sizeof *tmp);
        free(tmp);
        napi_set_named_property(env, out, "stats", v);
    }
    if ((o->cfg.flags & APM_F_KEEP_FS) && (o->cfg.flags & APM_F_ZSCORE) && o->cfg.num_lags > 0) {
        const uint64_t cap = (uint64_t)ticks * o->keys * o->cfg.num_lags;
        const bool bound = o->fs_pinned && cap <= o->fs_pinned_cap;  /* the records are already there */
        apm_fs_record* tmp = bound ? o->fs_pinned : (apm_fs_record*)malloc((cap ? cap : 1) * sizeof *tmp);
        apm_status st = apm_fetch_fs(o->ctx, tmp, bound ? o->fs_pinned_cap : cap, &n);
        if (st) { if (!bound) free(tmp); return fail(env, o->ctx, st, "apm_fetch_fs"); }
        napi_create_arraybuffer(env, n * sizeof *tmp, &p, &v);
        memcpy(p, tmp, n * sizeof *tmp);
        if (!bound) free(tmp);
        napi_set_named_property(env, out, "fs", v);
        if (o->cfg.flags & APM_F_ALERTS) {
            const uint64_t acap = (uint64_t)ticks * o->cfg.max_alerts_per_tick;
            tmp = (apm_fs_record*)malloc((acap ? acap : 1) * sizeof *tmp);
            st = apm_fetch_alerts(o->ctx, tmp, acap, &n);
            if (st) { free(tmp); return fail(env, o->ctx, st, "apm_fetch_alerts"); }
            napi_create_arraybuffer(env, n * sizeof *tmp, &p, &v);
            memcpy(p, tmp, n * sizeof *tmp);
            free(tmp);
            napi_set_named_property(env, out, "alerts", v);
        }
    }
    if (o->comm_ranks > 0) {
        /* multi-GPU: the alert records of ALL ranks for the same ticks (every rank calls at the same point of its stream) */
        uint64_t cap = o->gather_cap;
        if (!cap) cap = (uint64_t)o->comm_ranks * o->cfg.max_alerts_per_tick * (ticks ? ticks : 1);
        apm_status st = gather_buffer(o, cap);
        if (st) return fail(env, NULL, st, "apm_host_alloc");
        st = apm_gather_alerts(o->ctx, o->gather_pinned, o->gather_cap, &n);
        if (st) return fail(env, o->ctx, st, "apm_gather_alerts");
        napi_create_arraybuffer(env, n * sizeof(apm_fs_record), &p, &v);
        memcpy(p, o->gather_pinned, n * sizeof(apm_fs_record));
        napi_set_named_property(env, out, "gathered", v);
    }
    if (o->cfg.flags & APM_F_PASSTHROUGH) {
        /* writeHeapToQueue's output of the retained ticks (stream_calc_stats.js:136-155): endTs order, arrival sequence numbers */
        const uint64_t cap = 2 * o->cfg.max_tx_per_bucket * (uint64_t)(o->cfg.buffer_buckets + 3);
        int64_t* ts = (int64_t*)malloc((cap ? cap : 1) * sizeof *ts);
        uint32_t* seq = (uint32_t*)malloc((cap ? cap : 1) * sizeof *seq);
        uint32_t* counts = (uint32_t*)calloc(ticks ? ticks : 1, sizeof *counts);
        uint32_t nt = 0;
        apm_status st = apm_fetch_passthrough(o->ctx, ts, seq, cap, &n, counts, ticks, &nt);
        if (st) { free(ts); free(seq); free(counts); return fail(env, o->ctx, st, "apm_fetch_passthrough"); }
        napi_value pt;
        napi_create_object(env, &pt);
        napi_create_arraybuffer(env, n * sizeof *ts, &p, &v); memcpy(p, ts, n * sizeof *ts); napi_set_named_property(env, pt, "endTs", v);
        napi_create_arraybuffer(env, n * sizeof *seq, &p, &v); memcpy(p, seq, n * sizeof *seq); napi_set_named_property(env, pt, "seq", v);
        napi_create_arraybuffer(env, nt * sizeof *counts, &p, &v); memcpy(p, counts, nt * sizeof *counts); napi_set_named_property(env, pt, "tickCounts", v);
        free(ts); free(seq); free(counts);
        napi_set_named_property(env, out, "passthrough", pt);
    }
    apm_status st = apm_release_ticks(o->ctx);
    if (st) return fail(env, o->ctx, st, "apm_release_ticks");
    return out;
}

/* tick(latestBucket) */
static napi_value Tick(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    int64_t latest = 0;
    if (argc < 1 || napi_get_value_int64(env, argv[0], &latest) != napi_ok) {
        napi_throw_error(env, "APM_B200", "tick(latestBucket): number expected");
        return NULL;
    }
    apm_tick_result res;
    apm_status st = apm_tick(o->ctx, latest, &res);
    if (st) return fail(env, o->ctx, st, "apm_tick");
    return collect(env, o, &res);
}

/* processColumns(svc, jvm, endTs, elapsed): a run of consumeMsg calls, ticks found on the device */
static napi_value ProcessColumns(napi_env env, napi_callback_info info) {
    size_t argc = 4, n = 0;
    napi_value argv[4];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    apm_tx_columns src;
    if (argc < 4 || !get_columns(env, argv, &src, &n)) {
        napi_throw_error(env, "APM_B200", "processColumns(Uint32Array, Uint32Array, BigInt64Array, Int32Array) of equal length");
        return NULL;
    }
    apm_tick_result res;
    apm_status st = apm_process_host(o->ctx, &src, n, &res);
    if (st) return fail(env, o->ctx, st, "apm_process_host");
    return collect(env, o, &res);
}

/* setKeyParams(key, lagIndex, threshold, influence) */
/* processPacked(key, endTsDelta, elapsed: Uint32Array x3, baseTsMs): 12-byte transactions, see apm_process_host_packed */
static napi_value ProcessPacked(napi_env env, napi_callback_info info) {
    size_t argc = 4;
    napi_value argv[4];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    napi_typedarray_type t[3];
    size_t len[3] = {0, 0, 0};
    void* p[3] = {NULL, NULL, NULL};
    int64_t base = 0;
    bool ok = argc >= 4 && napi_get_value_int64(env, argv[3], &base) == napi_ok;
    for (int i = 0; ok && i < 3; ++i)
        ok = napi_get_typedarray_info(env, argv[i], &t[i], &len[i], &p[i], NULL, NULL) == napi_ok && t[i] == napi_uint32_array;
    if (!ok || len[0] != len[1] || len[0] != len[2]) {
        napi_throw_error(env, "APM_B200", "processPacked(Uint32Array, Uint32Array, Uint32Array, baseTsMs) of equal length");
        return NULL;
    }
    const apm_tx_packed src = {(const uint32_t*)p[0], (const uint32_t*)p[1], (const uint32_t*)p[2]};
    apm_tick_result res;
    apm_status st = apm_process_host_packed(o->ctx, &src, base, len[0], &res);
    if (st) return fail(env, o->ctx, st, "apm_process_host_packed");
    return collect(env, o, &res);
}

static napi_value SetKeyParams(napi_env env, napi_callback_info info) {
    size_t argc = 4;
    napi_value argv[4];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    double v[4];
    for (int i = 0; i < 4; ++i)
        if (argc < 4 || napi_get_value_double(env, argv[i], &v[i]) != napi_ok) {
            napi_throw_error(env, "APM_B200", "setKeyParams(key, lagIndex, threshold, influence): numbers expected");
            return NULL;
        }
    apm_status st = apm_set_key_params(o->ctx, (uint32_t)v[0], 1, (int32_t)v[1], &v[2], &v[3]);
    if (st) return fail(env, o->ctx, st, "apm_set_key_params");
    return undefined(env);
}

static napi_value state_io(napi_env env, napi_callback_info info, int save) {
    size_t argc = 1, len = 0;
    napi_value argv[1];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    char path[1024];
    if (argc < 1 || napi_get_value_string_utf8(env, argv[0], path, sizeof path, &len) != napi_ok) {
        napi_throw_error(env, "APM_B200", "path string expected");
        return NULL;
    }
    apm_status st = save ? apm_state_save(o->ctx, path) : apm_state_load(o->ctx, path);
    if (st) return fail(env, o->ctx, st, save ? "apm_state_save" : "apm_state_load");
    return undefined(env);
}
static napi_value SaveState(napi_env env, napi_callback_info info) { return state_io(env, info, 1); }
static napi_value LoadState(napi_env env, napi_callback_info info) { return state_io(env, info, 0); }

/* ---- log feeder: stream_parse_transactions.js:733-791 readLine over a whole buffer ---- */

/* dictSet(namesArray, idsUint32Array): rewritten service names -> service ids of the device parser */
static napi_value DictSet(napi_env env, napi_callback_info info) {
    size_t argc = 2;
    napi_value argv[2];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    bool is_arr = false;
    uint32_t n = 0;
    napi_typedarray_type t;
    size_t len = 0;
    void* ids = NULL;
    if (argc < 2 || napi_is_array(env, argv[0], &is_arr) != napi_ok || !is_arr || napi_get_array_length(env, argv[0], &n) != napi_ok ||
        napi_get_typedarray_info(env, argv[1], &t, &len, &ids, NULL, NULL) != napi_ok || t != napi_uint32_array || len != n) {
        napi_throw_error(env, "APM_B200", "dictSet(names: string[], ids: Uint32Array) of equal length");
        return NULL;
    }
    uint64_t* hashes = (uint64_t*)malloc((n ? n : 1) * sizeof *hashes);
    char name[1024];
    for (uint32_t i = 0; i < n; ++i) {
        napi_value e;
        size_t got = 0;
        if (napi_get_element(env, argv[0], i, &e) != napi_ok || napi_get_value_string_utf8(env, e, name, sizeof name, &got) != napi_ok) {
            free(hashes);
            napi_throw_error(env, "APM_B200", "dictSet: names must be strings shorter than 1024 bytes");
            return NULL;
        }
        hashes[i] = apm_name_hash(name, got);
    }
    apm_status st = apm_dict_set(o->ctx, hashes, (const uint32_t*)ids, n);
    free(hashes);
    if (st) return fail(env, o->ctx, st, "apm_dict_set");
    return undefined(env);
}

/* parseLog(text: Uint8Array, fileType, jvmId[, tz: number[] = [localFromMs, offsetMs, ...]]) ->
 * { records: ArrayBuffer of 64-byte apm_log_record, lines, exits, fed, unknownService, malformed } */
static napi_value ParseLog(napi_env env, napi_callback_info info) {
    size_t argc = 4;
    napi_value argv[4];
    addon_obj* o = self_of(env, info, &argc, argv);
    if (!o) return NULL;
    napi_typedarray_type t;
    size_t len = 0;
    void* text = NULL;
    uint32_t file_type = 0, jvm = 0;
    if (argc < 3 || napi_get_typedarray_info(env, argv[0], &t, &len, &text, NULL, NULL) != napi_ok || t != napi_uint8_array ||
        napi_get_value_uint32(env, argv[1], &file_type) != napi_ok || napi_get_value_uint32(env, argv[2], &jvm) != napi_ok) {
        napi_throw_error(env, "APM_B200", "parseLog(text: Uint8Array, fileType: 0|1, jvmId[, tzRules: number[]])");
        return NULL;
    }
    apm_tz_rule tz[APM_MAX_TZ_RULES];
    uint32_t ntz = 0;
    bool is_arr = false;
    if (argc >= 4 && napi_is_array(env, argv[3], &is_arr) == napi_ok && is_arr) {
        uint32_t m = 0;
        napi_get_array_length(env, argv[3], &m);
        if (m % 2 || m / 2 > APM_MAX_TZ_RULES) {
            napi_throw_error(env, "APM_B200", "parseLog: tzRules is a flat array of (localFromMs, offsetMs) pairs, at most 8");
            return NULL;
        }
        for (uint32_t i = 0; i < m; ++i) {
            napi_value e;
            double d = 0;
            napi_get_element(env, argv[3], i, &e);
            if (napi_get_value_double(env, e, &d) != napi_ok) { napi_throw_error(env, "APM_B200", "parseLog: tzRules must be numbers"); return NULL; }
            if (i % 2) tz[i / 2].offset_ms = (int64_t)d; else tz[i / 2].local_from_ms = d < -9e18 ? INT64_MIN : (int64_t)d;
        }
        ntz = m / 2;
    }
    uint64_t cap = 1;
    for (size_t i = 0; i < len; ++i) cap += ((const unsigned char*)text)[i] == '\n';
    apm_log_record* tmp = (apm_log_record*)malloc(cap * sizeof *tmp);
    uint64_t n = 0;
    apm_parse_result pr;
    apm_status st = apm_parse_log_host(o->ctx, (const char*)text, len, file_type, jvm, ntz ? tz : NULL, ntz, tmp, cap, &n, &pr);
    if (st) { free(tmp); return fail(env, o->ctx, st, "apm_parse_log_host"); }
    napi_value out, v;
    void* p;
    napi_create_object(env, &out);
    napi_create_arraybuffer(env, n * sizeof *tmp, &p, &v);
    memcpy(p, tmp, n * sizeof *tmp);
    free(tmp);
    napi_set_named_property(env, out, "records", v);
    const struct { const char* name; uint64_t val; } f[] = {{"lines", pr.lines}, {"exits", pr.exits}, {"fed", pr.fed},
                                                            {"unknownService", pr.unknown_service}, {"malformed", pr.malformed}};
    for (size_t i = 0; i < sizeof f / sizeof f[0]; ++i) {
        napi_create_double(env, (double)f[i].val, &v);
        napi_set_named_property(env, out, f[i].name, v);
    }
    return out;
}

/* feedParsed(): the exits of the last parseLog go through stats + z-score on the device */
static napi_value FeedParsed(napi_env env, napi_callback_info info) {
    size_t argc = 0;
    addon_obj* o = self_of(env, info, &argc, NULL);
    if (!o) return NULL;
    apm_tick_result res;
    apm_status st = apm_feed_parsed(o->ctx, &res);
    if (st) return fail(env, o->ctx, st, "apm_feed_parsed");
    return collect(env, o, &res);
}


/* ---- the rest of the C ABI, method by method -------------------------------------------------------------- */
#define ARGS(nmax)                                   \
    size_t argc = (nmax);                            \
    napi_value argv[(nmax) ? (nmax) : 1];            \
    addon_obj* o = self_of(env, info, &argc, argv);  \
    if (!o) return NULL
#define BAD(msg) do { napi_throw_error(env, "APM_B200", msg); return NULL; } while (0)

static bool typed(napi_env env, napi_value v, napi_typedarray_type want, void** data, size_t* len) {
    napi_typedarray_type t;
    return napi_get_typedarray_info(env, v, &t, len, data, NULL, NULL) == napi_ok && t == want;
}
static napi_value number_of(napi_env env, double d) { napi_value v; napi_create_double(env, d, &v); return v; }
static napi_value buffer_of(napi_env env, const void* src, size_t bytes) {
    napi_value v;
    void* p;
    napi_create_arraybuffer(env, bytes, &p, &v);
    if (bytes) memcpy(p, src, bytes);
    return v;
}

/* dropLag(lagIndex): removeStaleLagData for one lag (stream_calc_z_score.js:170-193) */
static napi_value DropLag(napi_env env, napi_callback_info info) {
    ARGS(1);
    double l;
    if (argc < 1 || napi_get_value_double(env, argv[0], &l) != napi_ok) BAD("dropLag(lagIndex)");
    apm_status st = apm_drop_lag(o->ctx, (int32_t)l);
    if (st) return fail(env, o->ctx, st, "apm_drop_lag");
    return undefined(env);
}

/* setAlertRules({enabled, window, required, alertOnBothOnly, suppressedLagMask, hardMaxMs, hardMinMs, hardMinTpm}):
 * the counter core of processFSEntry (stream_process_alerts.js:348-434); defaults config/apm_config.json:189-204 */
static napi_value SetAlertRules(napi_env env, napi_callback_info info) {
    ARGS(1);
    if (argc < 1) BAD("setAlertRules(options)");
    apm_alert_rules r;
    double d;
    memset(&r, 0, sizeof r);
    r.struct_size = sizeof r;
    r.enabled = get_number(env, argv[0], "enabled", &d) ? (int32_t)d : 1;
    r.window = get_number(env, argv[0], "window", &d) ? (int32_t)d : 60;
    r.required = get_number(env, argv[0], "required", &d) ? (int32_t)d : 45;
    r.alert_on_both_only = get_number(env, argv[0], "alertOnBothOnly", &d) ? (int32_t)d : 1;
    r.suppressed_lag_mask = get_number(env, argv[0], "suppressedLagMask", &d) ? (uint32_t)d : 0;
    r.hard_max_ms = get_number(env, argv[0], "hardMaxMs", &d) ? d : 10000.0;
    r.hard_min_ms = get_number(env, argv[0], "hardMinMs", &d) ? d : 200.0;
    r.hard_min_tpm = get_number(env, argv[0], "hardMinTpm", &d) ? d : 1.0;
    apm_status st = apm_set_alert_rules(o->ctx, &r);
    if (st) return fail(env, o->ctx, st, "apm_set_alert_rules");
    return undefined(env);
}

/* setAlertKeyParams(keyLo, hardMaxMs: Float64Array, suppressed: Uint8Array) */
static napi_value SetAlertKeyParams(napi_env env, napi_callback_info info) {
    ARGS(3);
    double lo;
    void *h, *s;
    size_t nh, ns;
    if (argc < 3 || napi_get_value_double(env, argv[0], &lo) != napi_ok || !typed(env, argv[1], napi_float64_array, &h, &nh) ||
        !typed(env, argv[2], napi_uint8_array, &s, &ns) || nh != ns)
        BAD("setAlertKeyParams(keyLo, hardMaxMs: Float64Array, suppressed: Uint8Array) of equal length");
    apm_status st = apm_set_alert_key_params(o->ctx, (uint32_t)lo, (uint32_t)nh, (const double*)h, (const uint8_t*)s);
    if (st) return fail(env, o->ctx, st, "apm_set_alert_key_params");
    return undefined(env);
}

/* setKeyParamsRange(keyLo, lagIndex, thresholds: Float64Array, influences: Float64Array): updateAllServiceSettings in one call */
static napi_value SetKeyParamsRange(napi_env env, napi_callback_info info) {
    ARGS(4);
    double lo, l;
    void *t, *f;
    size_t nt, nf;
    if (argc < 4 || napi_get_value_double(env, argv[0], &lo) != napi_ok || napi_get_value_double(env, argv[1], &l) != napi_ok ||
        !typed(env, argv[2], napi_float64_array, &t, &nt) || !typed(env, argv[3], napi_float64_array, &f, &nf) || nt != nf)
        BAD("setKeyParamsRange(keyLo, lagIndex, thresholds: Float64Array, influences: Float64Array) of equal length");
    apm_status st = apm_set_key_params(o->ctx, (uint32_t)lo, (uint32_t)nt, (int32_t)l, (const double*)t, (const double*)f);
    if (st) return fail(env, o->ctx, st, "apm_set_key_params");
    return undefined(env);
}

/* ringLoad(key, lagIndex, series, values: Float64Array): one avgList / per75List / per95List of a resume file (NaN = null) */
static napi_value RingLoad(napi_env env, napi_callback_info info) {
    ARGS(4);
    double k, l, s;
    void* v;
    size_t n;
    if (argc < 4 || napi_get_value_double(env, argv[0], &k) != napi_ok || napi_get_value_double(env, argv[1], &l) != napi_ok ||
        napi_get_value_double(env, argv[2], &s) != napi_ok || !typed(env, argv[3], napi_float64_array, &v, &n))
        BAD("ringLoad(key, lagIndex, series, values: Float64Array)");
    apm_status st = apm_ring_load(o->ctx, (uint32_t)k, (int32_t)l, (int32_t)s, (const double*)v, (uint32_t)n);
    if (st) return fail(env, o->ctx, st, "apm_ring_load");
    return undefined(env);
}

/* ringLoadAll(lagIndex, values: Float64Array [len][3][keys], len) */
static napi_value RingLoadAll(napi_env env, napi_callback_info info) {
    ARGS(3);
    double l, len;
    void* v;
    size_t n;
    if (argc < 3 || napi_get_value_double(env, argv[0], &l) != napi_ok || !typed(env, argv[1], napi_float64_array, &v, &n) ||
        napi_get_value_double(env, argv[2], &len) != napi_ok || n != (size_t)len * 3 * o->keys)
        BAD("ringLoadAll(lagIndex, values: Float64Array of len * 3 * keys, len)");
    apm_status st = apm_ring_load_all(o->ctx, (int32_t)l, (const double*)v, (uint32_t)len, 0);
    if (st) return fail(env, o->ctx, st, "apm_ring_load_all");
    return undefined(env);
}

/* ringRead(key, lagIndex, series) -> ArrayBuffer of doubles, oldest first */
static napi_value RingRead(napi_env env, napi_callback_info info) {
    ARGS(3);
    double k, l, s;
    if (argc < 3 || napi_get_value_double(env, argv[0], &k) != napi_ok || napi_get_value_double(env, argv[1], &l) != napi_ok ||
        napi_get_value_double(env, argv[2], &s) != napi_ok || (int)l < 0 || (int)l >= o->cfg.num_lags)
        BAD("ringRead(key, lagIndex, series)");
    const uint32_t cap = (uint32_t)o->cfg.lags[(int)l];
    double* tmp = (double*)malloc((size_t)cap * sizeof *tmp);
    uint32_t len = 0;
    apm_status st = apm_ring_read(o->ctx, (uint32_t)k, (int32_t)l, (int32_t)s, tmp, cap, &len);
    if (st) { free(tmp); return fail(env, o->ctx, st, "apm_ring_read"); }
    napi_value v = buffer_of(env, tmp, (size_t)len * sizeof *tmp);
    free(tmp);
    return v;
}

/* setLatest(label) / markSeen(keyLo, n) / logRead(label) / latest(): interop with the reference's JSON resume files */
static napi_value SetLatest(napi_env env, napi_callback_info info) {
    ARGS(1);
    int64_t lab;
    if (argc < 1 || napi_get_value_int64(env, argv[0], &lab) != napi_ok) BAD("setLatest(latestBucket)");
    apm_status st = apm_set_latest(o->ctx, lab);
    if (st) return fail(env, o->ctx, st, "apm_set_latest");
    return undefined(env);
}
static napi_value MarkSeen(napi_env env, napi_callback_info info) {
    ARGS(2);
    double lo, n;
    if (argc < 2 || napi_get_value_double(env, argv[0], &lo) != napi_ok || napi_get_value_double(env, argv[1], &n) != napi_ok) BAD("markSeen(keyLo, n)");
    apm_status st = apm_mark_seen(o->ctx, (uint32_t)lo, (uint32_t)n);
    if (st) return fail(env, o->ctx, st, "apm_mark_seen");
    return undefined(env);
}
static napi_value LogRead(napi_env env, napi_callback_info info) {
    ARGS(1);
    int64_t lab;
    if (argc < 1 || napi_get_value_int64(env, argv[0], &lab) != napi_ok) BAD("logRead(label)");
    uint64_t n = 0;
    apm_status st = apm_log_read(o->ctx, lab, NULL, 0, &n);
    if (st) return fail(env, o->ctx, st, "apm_log_read");
    uint64_t* tmp = (uint64_t*)malloc((n ? n : 1) * sizeof *tmp);
    if (n) st = apm_log_read(o->ctx, lab, tmp, n, &n);
    if (st) { free(tmp); return fail(env, o->ctx, st, "apm_log_read"); }
    napi_value v = buffer_of(env, tmp, n * sizeof *tmp);
    free(tmp);
    return v;
}
static napi_value Latest(napi_env env, napi_callback_info info) {
    ARGS(0);
    napi_value v;
    napi_create_int64(env, apm_latest(o->ctx), &v);
    return v;
}

/* setTickMarks(marks: BigInt64Array of (pos, label) pairs): the next process* call ticks where the host says */
static napi_value SetTickMarks(napi_env env, napi_callback_info info) {
    ARGS(1);
    void* p;
    size_t n;
    if (argc < 1 || !typed(env, argv[0], napi_bigint64_array, &p, &n) || n % 2) BAD("setTickMarks(BigInt64Array of (pos, label) pairs)");
    apm_status st = apm_set_tick_marks(o->ctx, (const apm_tick_mark*)p, (uint32_t)(n / 2));
    if (st) return fail(env, o->ctx, st, "apm_set_tick_marks");
    return undefined(env);
}

/* processPacked8(words: BigUint64Array, escapes: Int32Array of (index, elapsed) pairs, baseTsMs) */
static napi_value ProcessPacked8(napi_env env, napi_callback_info info) {
    ARGS(3);
    void *w, *e;
    size_t nw, ne;
    int64_t base;
    if (argc < 3 || !typed(env, argv[0], napi_biguint64_array, &w, &nw) || !typed(env, argv[1], napi_int32_array, &e, &ne) || ne % 2 ||
        napi_get_value_int64(env, argv[2], &base) != napi_ok)
        BAD("processPacked8(words: BigUint64Array, escapes: Int32Array of (index, elapsed) pairs, baseTsMs)");
    apm_tick_result res;
    apm_status st = apm_process_host_packed8(o->ctx, (const uint64_t*)w, (const apm_tx_escape*)e, (uint32_t)(ne / 2), base, nw, &res);
    if (st) return fail(env, o->ctx, st, "apm_process_host_packed8");
    return collect(env, o, &res);
}

/* bindFsOutput(): fs records stream into a page-locked buffer while the following ticks run */
static napi_value BindFsOutput(napi_env env, napi_callback_info info) {
    ARGS(0);
    const uint64_t cap = (uint64_t)o->cfg.out_ticks * o->keys * (uint64_t)(o->cfg.num_lags > 0 ? o->cfg.num_lags : 1);
    if (!o->fs_pinned) {
        void* p = NULL;
        apm_status st = apm_host_alloc(cap * sizeof(apm_fs_record), &p);
        if (st) return fail(env, NULL, st, "apm_host_alloc");
        o->fs_pinned = (apm_fs_record*)p;
        o->fs_pinned_cap = cap;
    }
    apm_status st = apm_bind_fs_output(o->ctx, o->fs_pinned, o->fs_pinned_cap);
    if (st) return fail(env, o->ctx, st, "apm_bind_fs_output");
    return undefined(env);
}

/* ---- multi-GPU: commInit(rank, nranks, id: Uint8Array(128)), commSetCapacity(records), bindGatherOutput(capRecords),
 *      gatherAlerts(capRecords) -> ArrayBuffer of 128-byte records (sender's rank in pad_[2]) */
static napi_value CommInit(napi_env env, napi_callback_info info) {
    ARGS(3);
    double rank, nranks;
    void* id;
    size_t n;
    if (argc < 3 || napi_get_value_double(env, argv[0], &rank) != napi_ok || napi_get_value_double(env, argv[1], &nranks) != napi_ok ||
        !typed(env, argv[2], napi_uint8_array, &id, &n) || n != 128)
        BAD("commInit(rank, nranks, uniqueId: Uint8Array(128))");
    apm_status st = apm_comm_init(o->ctx, (int32_t)rank, (int32_t)nranks, id);
    if (st) return fail(env, o->ctx, st, "apm_comm_init");
    o->comm_ranks = (int32_t)nranks;
    return undefined(env);
}
static napi_value CommSetCapacity(napi_env env, napi_callback_info info) {
    ARGS(1);
    double n;
    if (argc < 1 || napi_get_value_double(env, argv[0], &n) != napi_ok) BAD("commSetCapacity(recordsPerBlock)");
    apm_status st = apm_comm_set_capacity(o->ctx, (uint32_t)n);
    if (st) return fail(env, o->ctx, st, "apm_comm_set_capacity");
    return undefined(env);
}
static apm_status gather_buffer(addon_obj* o, uint64_t cap) {
    if (o->gather_pinned && o->gather_cap >= cap) return APM_OK;
    apm_bind_gather_output(o->ctx, NULL, 0);
    apm_host_free(o->gather_pinned);
    o->gather_pinned = NULL;
    void* p = NULL;
    apm_status st = apm_host_alloc((cap ? cap : 1) * sizeof(apm_fs_record), &p);
    if (st) return st;
    o->gather_pinned = (apm_fs_record*)p;
    o->gather_cap = cap;
    return APM_OK;
}
static napi_value BindGatherOutput(napi_env env, napi_callback_info info) {
    ARGS(1);
    double cap;
    if (argc < 1 || napi_get_value_double(env, argv[0], &cap) != napi_ok) BAD("bindGatherOutput(capRecords)");
    apm_status st = gather_buffer(o, (uint64_t)cap);
    if (st) return fail(env, NULL, st, "apm_host_alloc");
    st = apm_bind_gather_output(o->ctx, o->gather_pinned, o->gather_cap);
    if (st) return fail(env, o->ctx, st, "apm_bind_gather_output");
    return undefined(env);
}
static napi_value GatherAlerts(napi_env env, napi_callback_info info) {
    ARGS(1);
    double cap;
    if (argc < 1 || napi_get_value_double(env, argv[0], &cap) != napi_ok) BAD("gatherAlerts(capRecords)");
    apm_status st = gather_buffer(o, (uint64_t)cap);
    if (st) return fail(env, NULL, st, "apm_host_alloc");
    uint64_t n = 0;
    st = apm_gather_alerts(o->ctx, o->gather_pinned, o->gather_cap, &n);
    if (st) return fail(env, o->ctx, st, "apm_gather_alerts");
    return buffer_of(env, o->gather_pinned, n * sizeof(apm_fs_record));
}

/* counters() -> { txIngested, txDropped, txInvalid, ticks, statRecords, fsRecords, alerts, h2dBytes, d2hBytes, kernelLaunches, keysSeen, incFallbacks } */
static napi_value Counters(napi_env env, napi_callback_info info) {
    ARGS(0);
    apm_counters c;
    apm_status st = apm_get_counters(o->ctx, &c);
    if (st) return fail(env, o->ctx, st, "apm_get_counters");
    napi_value out;
    napi_create_object(env, &out);
    const struct { const char* name; uint64_t v; } f[] = {
        {"txIngested", c.tx_ingested}, {"txDropped", c.tx_dropped}, {"txInvalid", c.tx_invalid}, {"ticks", c.ticks},
        {"statRecords", c.stat_records}, {"fsRecords", c.fs_records}, {"alerts", c.alerts}, {"h2dBytes", c.h2d_bytes},
        {"d2hBytes", c.d2h_bytes}, {"kernelLaunches", c.kernel_launches}, {"keysSeen", c.keys_seen}, {"incFallbacks", c.inc_fallbacks}};
    for (size_t i = 0; i < sizeof f / sizeof f[0]; ++i) napi_set_named_property(env, out, f[i].name, number_of(env, (double)f[i].v));
    return out;
}
static napi_value ResetCounters(napi_env env, napi_callback_info info) {
    ARGS(0);
    apm_status st = apm_reset_counters(o->ctx);
    if (st) return fail(env, o->ctx, st, "apm_reset_counters");
    return undefined(env);
}
static napi_value Sync(napi_env env, napi_callback_info info) {
    ARGS(0);
    apm_status st = apm_sync(o->ctx);
    if (st) return fail(env, o->ctx, st, "apm_sync");
    return undefined(env);
}
static napi_value TicksBuffered(napi_env env, napi_callback_info info) {
    ARGS(0);
    return number_of(env, (double)apm_ticks_buffered(o->ctx));
}

/* measurement helpers (no reference counterpart): profileEnable(on), profileRead() -> { ms: ArrayBuffer f64[12], spans: ArrayBuffer u64[12] },
 * setOverlap(on), timerStart(), timerStop() -> ms */
static napi_value ProfileEnable(napi_env env, napi_callback_info info) {
    ARGS(1);
    double on = 1;
    if (argc >= 1) napi_get_value_double(env, argv[0], &on);
    apm_status st = apm_profile_enable(o->ctx, on != 0);
    if (st) return fail(env, o->ctx, st, "apm_profile_enable");
    return undefined(env);
}
static napi_value ProfileRead(napi_env env, napi_callback_info info) {
    ARGS(0);
    apm_kernel_times kt;
    apm_status st = apm_profile_read(o->ctx, &kt);
    if (st) return fail(env, o->ctx, st, "apm_profile_read");
    napi_value out;
    napi_create_object(env, &out);
    napi_set_named_property(env, out, "ms", buffer_of(env, kt.ms, sizeof kt.ms));
    napi_set_named_property(env, out, "spans", buffer_of(env, kt.spans, sizeof kt.spans));
    return out;
}
static napi_value SetOverlap(napi_env env, napi_callback_info info) {
    ARGS(1);
    double on = 1;
    if (argc >= 1) napi_get_value_double(env, argv[0], &on);
    apm_status st = apm_set_overlap(o->ctx, on != 0);
    if (st) return fail(env, o->ctx, st, "apm_set_overlap");
    return undefined(env);
}
static napi_value TimerStart(napi_env env, napi_callback_info info) {
    ARGS(0);
    apm_status st = apm_timer_start(o->ctx);
    if (st) return fail(env, o->ctx, st, "apm_timer_start");
    return undefined(env);
}
static napi_value TimerStop(napi_env env, napi_callback_info info) {
    ARGS(0);
    double ms = 0;
    apm_status st = apm_timer_stop(o->ctx, &ms);
    if (st) return fail(env, o->ctx, st, "apm_timer_stop");
    return number_of(env, ms);
}

/* ---- module-level functions (no context) ------------------------------------------------------------------- */
static size_t plain_args(napi_env env, napi_callback_info info, size_t want, napi_value* argv) {
    size_t argc = want;
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    return argc;
}
/* quantize(x, digits) = parseFloat(x.toFixed(digits)) as the device computes it (entries.js:65-73 -> 58-61) */
static napi_value Quantize(napi_env env, napi_callback_info info) {
    napi_value argv[2];
    double x, d;
    if (plain_args(env, info, 2, argv) < 2 || napi_get_value_double(env, argv[0], &x) != napi_ok || napi_get_value_double(env, argv[1], &d) != napi_ok)
        BAD("quantize(x, digits)");
    return number_of(env, apm_quantize(x, (int32_t)d));
}
static napi_value OwnerOfKey(napi_env env, napi_callback_info info) {
    napi_value argv[2];
    double k, n;
    if (plain_args(env, info, 2, argv) < 2 || napi_get_value_double(env, argv[0], &k) != napi_ok || napi_get_value_double(env, argv[1], &n) != napi_ok)
        BAD("ownerOfKey(key, nranks)");
    return number_of(env, (double)apm_owner_of_key((uint64_t)k, (uint32_t)n));
}
/* ownerTableBuild(counts: BigUint64Array, nranks, keyCost) -> ArrayBuffer of uint32 owners */
static napi_value OwnerTableBuild(napi_env env, napi_callback_info info) {
    napi_value argv[3];
    void* c;
    size_t n;
    double nranks, cost;
    if (plain_args(env, info, 3, argv) < 3 || !typed(env, argv[0], napi_biguint64_array, &c, &n) || napi_get_value_double(env, argv[1], &nranks) != napi_ok ||
        napi_get_value_double(env, argv[2], &cost) != napi_ok)
        BAD("ownerTableBuild(counts: BigUint64Array, nranks, keyCost)");
    uint32_t* owner = (uint32_t*)malloc((n ? n : 1) * sizeof *owner);
    apm_status st = apm_owner_table_build((const uint64_t*)c, (uint32_t)n, (uint32_t)nranks, (uint64_t)cost, owner);
    if (st) { free(owner); return fail(env, NULL, st, "apm_owner_table_build"); }
    napi_value v = buffer_of(env, owner, n * sizeof *owner);
    free(owner);
    return v;
}
/* ownerTableBuildCapped(counts: BigUint64Array, nranks, keySlack) -> ArrayBuffer of uint32 owners */
static napi_value OwnerTableBuildCapped(napi_env env, napi_callback_info info) {
    napi_value argv[3];
    void* c;
    size_t n;
    double nranks, slack;
    if (plain_args(env, info, 3, argv) < 3 || !typed(env, argv[0], napi_biguint64_array, &c, &n) || napi_get_value_double(env, argv[1], &nranks) != napi_ok ||
        napi_get_value_double(env, argv[2], &slack) != napi_ok)
        BAD("ownerTableBuildCapped(counts: BigUint64Array, nranks, keySlack)");
    uint32_t* owner = (uint32_t*)malloc((n ? n : 1) * sizeof *owner);
    apm_status st = apm_owner_table_build_capped((const uint64_t*)c, (uint32_t)n, (uint32_t)nranks, slack, owner);
    if (st) { free(owner); return fail(env, NULL, st, "apm_owner_table_build_capped"); }
    napi_value v = buffer_of(env, owner, n * sizeof *owner);
    free(owner);
    return v;
}
/* commUniqueId() -> ArrayBuffer(128) for commInit on every rank */
static napi_value CommUniqueId(napi_env env, napi_callback_info info) {
    (void)info;
    unsigned char id[128];
    apm_status st = apm_comm_unique_id(id);
    if (st) return fail(env, NULL, st, "apm_comm_unique_id");
    return buffer_of(env, id, sizeof id);
}
/* preloadKernels(device) */
static napi_value PreloadKernels(napi_env env, napi_callback_info info) {
    napi_value argv[1];
    double dev = 0;
    if (plain_args(env, info, 1, argv) >= 1) napi_get_value_double(env, argv[0], &dev);
    apm_status st = apm_preload_kernels((int32_t)dev);
    if (st) return fail(env, NULL, st, "apm_preload_kernels");
    return undefined(env);
}
/* nameHash(name) -> the dictionary hash of a rewritten service name, as a decimal string of the u64 */
static napi_value NameHash(napi_env env, napi_callback_info info) {
    napi_value argv[1], v;
    char name[1024], out[32];
    size_t got = 0;
    if (plain_args(env, info, 1, argv) < 1 || napi_get_value_string_utf8(env, argv[0], name, sizeof name, &got) != napi_ok) BAD("nameHash(name)");
    snprintf(out, sizeof out, "%llu", (unsigned long long)apm_name_hash(name, got));
    napi_create_string_utf8(env, out, NAPI_AUTO_LENGTH, &v);
    return v;
}
/* calcPercentiles(values: Int32Array, percentiles: number[], device = 0[, noAlloc]) -> { sorted: ArrayBuffer of int32, pct: ArrayBuffer of f64 }:
 * Array.prototype.binaryConcat + calcPercentile (util_methods.js:102-106, 112-142) for one array */
static napi_value CalcPercentiles(napi_env env, napi_callback_info info) {
    napi_value argv[3];
    void* vals;
    size_t n;
    double pcts[APM_MAX_PCT], out[APM_MAX_PCT], dev = 0;
    const size_t argc = plain_args(env, info, 3, argv);
    bool is_arr = false;
    uint32_t np = 0;
    if (argc < 2 || !typed(env, argv[0], napi_int32_array, &vals, &n) || napi_is_array(env, argv[1], &is_arr) != napi_ok || !is_arr ||
        napi_get_array_length(env, argv[1], &np) != napi_ok || np > APM_MAX_PCT)
        BAD("calcPercentiles(values: Int32Array, percentiles: number[] (at most 4)[, device])");
    for (uint32_t i = 0; i < np; ++i) {
        napi_value e;
        napi_get_element(env, argv[1], i, &e);
        if (napi_get_value_double(env, e, &pcts[i]) != napi_ok) BAD("calcPercentiles: percentiles must be numbers");
    }
    if (argc >= 3) napi_get_value_double(env, argv[2], &dev);
    int32_t* sorted = (int32_t*)malloc((n ? n : 1) * sizeof *sorted);
    apm_status st = apm_sorted_percentiles_host((int32_t)dev, (const int32_t*)vals, (uint32_t)n, pcts, (int32_t)np, out, sorted);
    if (st) { free(sorted); return fail(env, NULL, st, "apm_sorted_percentiles_host"); }
    napi_value res;
    napi_create_object(env, &res);
    napi_set_named_property(env, res, "sorted", buffer_of(env, sorted, n * sizeof *sorted));
    napi_set_named_property(env, res, "pct", buffer_of(env, out, np * sizeof(double)));
    free(sorted);
    return res;
}
/* partitionColumns(svc, jvm, endTs, elapsed, { bucketMs, latest, numServices, numJvms, nranks }[, ownerTable: Uint32Array[, localId: Uint32Array]])
 * -> { latest, n: ArrayBuffer u64[nranks], "0": { svc, jvm, endTs, elapsed, marks }, "1": ... }: host binning by key owner (apm_partition_host) */
static napi_value PartitionColumns(napi_env env, napi_callback_info info) {
    napi_value argv[7];
    const size_t argc = plain_args(env, info, 7, argv);
    apm_tx_columns src;
    size_t n = 0, nt = 0, nl = 0;
    double bucket, latest, S, J, R;
    void *table = NULL, *local = NULL;
    if (argc < 5 || !get_columns(env, argv, &src, &n) || !get_number(env, argv[4], "bucketMs", &bucket) || !get_number(env, argv[4], "numServices", &S) ||
        !get_number(env, argv[4], "numJvms", &J) || !get_number(env, argv[4], "nranks", &R) || R < 1)
        BAD("partitionColumns(svc, jvm, endTs, elapsed, { bucketMs, latest, numServices, numJvms, nranks }[, ownerTable[, localId]])");
    if (!get_number(env, argv[4], "latest", &latest)) latest = 0;
    if (argc >= 6 && !typed(env, argv[5], napi_uint32_array, &table, &nt)) table = NULL;
    if (argc >= 7 && !typed(env, argv[6], napi_uint32_array, &local, &nl)) local = NULL;
    const uint32_t nranks = (uint32_t)R;
    uint32_t marks_cap = 1;
    {   /* an upper bound on the ticks: distinct bucket labels in arrival order */
        int64_t run = (int64_t)latest;
        for (size_t i = 0; i < n; ++i) {
            const int64_t lab = src.end_ts_ms[i] >= 0 ? src.end_ts_ms[i] / (int64_t)bucket : -1;
            if (lab > run) { run = lab; marks_cap++; }
        }
    }
    apm_partition_out* outs = (apm_partition_out*)calloc(nranks, sizeof *outs);
    for (uint32_t r = 0; r < nranks; ++r) {
        outs[r].cols.service_id = (uint32_t*)malloc((n ? n : 1) * 4);
        outs[r].cols.jvm_id = (uint32_t*)malloc((n ? n : 1) * 4);
        outs[r].cols.end_ts_ms = (int64_t*)malloc((n ? n : 1) * 8);
        outs[r].cols.elapsed_ms = (int32_t*)malloc((n ? n : 1) * 4);
        outs[r].cap = n;
        outs[r].marks = (apm_tick_mark*)malloc(marks_cap * sizeof(apm_tick_mark));
        outs[r].marks_cap = marks_cap;
    }
    int64_t latest_out = 0;
    apm_status st = apm_partition_host(&src, n, (int64_t)bucket, (int64_t)latest, (uint32_t)S, (uint32_t)J, nranks, (const uint32_t*)table,
                                       (const uint32_t*)local, outs, &latest_out);
    napi_value res = NULL;
    if (!st) {
        napi_value v;
        napi_create_object(env, &res);
        napi_create_int64(env, latest_out, &v);
        napi_set_named_property(env, res, "latest", v);
        uint64_t* counts = (uint64_t*)calloc(nranks, sizeof *counts);
        for (uint32_t r = 0; r < nranks; ++r) {
            char name[16];
            napi_value part;
            napi_create_object(env, &part);
            counts[r] = outs[r].n;
            napi_set_named_property(env, part, "svc", buffer_of(env, outs[r].cols.service_id, outs[r].n * 4));
            napi_set_named_property(env, part, "jvm", buffer_of(env, outs[r].cols.jvm_id, outs[r].n * 4));
            napi_set_named_property(env, part, "endTs", buffer_of(env, outs[r].cols.end_ts_ms, outs[r].n * 8));
            napi_set_named_property(env, part, "elapsed", buffer_of(env, outs[r].cols.elapsed_ms, outs[r].n * 4));
            napi_set_named_property(env, part, "marks", buffer_of(env, outs[r].marks, outs[r].num_marks * sizeof(apm_tick_mark)));
            snprintf(name, sizeof name, "%u", r);
            napi_set_named_property(env, res, name, part);
        }
        napi_set_named_property(env, res, "n", buffer_of(env, counts, nranks * sizeof *counts));
        free(counts);
    }
    for (uint32_t r = 0; r < nranks; ++r) {
        free(outs[r].cols.service_id); free(outs[r].cols.jvm_id); free(outs[r].cols.end_ts_ms); free(outs[r].cols.elapsed_ms); free(outs[r].marks);
    }
    free(outs);
    if (st) return fail(env, NULL, st, "apm_partition_host");
    return res;
}

static napi_value Close(napi_env env, napi_callback_info info) {
    size_t argc = 0;
    napi_value self;
    addon_obj* o = NULL;
    napi_get_cb_info(env, info, &argc, NULL, &self, NULL);
    if (napi_unwrap(env, self, (void**)&o) == napi_ok && o && o->ctx) {
        apm_ctx_destroy(o->ctx);
        o->ctx = NULL;
        apm_host_free(o->fs_pinned); o->fs_pinned = NULL;
        apm_host_free(o->gather_pinned); o->gather_pinned = NULL;
    }
    return undefined(env);
}

napi_value napi_register_module_v1(napi_env env, napi_value exports) {
    const napi_property_descriptor methods[] = {
        {"addBatch", NULL, AddBatch, NULL, NULL, NULL, napi_default, NULL},
        {"tick", NULL, Tick, NULL, NULL, NULL, napi_default, NULL},
        {"processColumns", NULL, ProcessColumns, NULL, NULL, NULL, napi_default, NULL},
        {"processPacked", NULL, ProcessPacked, NULL, NULL, NULL, napi_default, NULL},
        {"setKeyParams", NULL, SetKeyParams, NULL, NULL, NULL, napi_default, NULL},
        {"saveState", NULL, SaveState, NULL, NULL, NULL, napi_default, NULL},
        {"loadState", NULL, LoadState, NULL, NULL, NULL, napi_default, NULL},
        {"dictSet", NULL, DictSet, NULL, NULL, NULL, napi_default, NULL},
        {"parseLog", NULL, ParseLog, NULL, NULL, NULL, napi_default, NULL},
        {"feedParsed", NULL, FeedParsed, NULL, NULL, NULL, napi_default, NULL},
        {"close", NULL, Close, NULL, NULL, NULL, napi_default, NULL},
        {"dropLag", NULL, DropLag, NULL, NULL, NULL, napi_default, NULL},
        {"setAlertRules", NULL, SetAlertRules, NULL, NULL, NULL, napi_default, NULL},
        {"setAlertKeyParams", NULL, SetAlertKeyParams, NULL, NULL, NULL, napi_default, NULL},
        {"setKeyParamsRange", NULL, SetKeyParamsRange, NULL, NULL, NULL, napi_default, NULL},
        {"ringLoad", NULL, RingLoad, NULL, NULL, NULL, napi_default, NULL},
        {"ringLoadAll", NULL, RingLoadAll, NULL, NULL, NULL, napi_default, NULL},
        {"ringRead", NULL, RingRead, NULL, NULL, NULL, napi_default, NULL},
        {"setLatest", NULL, SetLatest, NULL, NULL, NULL, napi_default, NULL},
        {"markSeen", NULL, MarkSeen, NULL, NULL, NULL, napi_default, NULL},
        {"logRead", NULL, LogRead, NULL, NULL, NULL, napi_default, NULL},
        {"latest", NULL, Latest, NULL, NULL, NULL, napi_default, NULL},
        {"setTickMarks", NULL, SetTickMarks, NULL, NULL, NULL, napi_default, NULL},
        {"processPacked8", NULL, ProcessPacked8, NULL, NULL, NULL, napi_default, NULL},
        {"bindFsOutput", NULL, BindFsOutput, NULL, NULL, NULL, napi_default, NULL},
        {"commInit", NULL, CommInit, NULL, NULL, NULL, napi_default, NULL},
        {"commSetCapacity", NULL, CommSetCapacity, NULL, NULL, NULL, napi_default, NULL},
        {"bindGatherOutput", NULL, BindGatherOutput, NULL, NULL, NULL, napi_default, NULL},
        {"gatherAlerts", NULL, GatherAlerts, NULL, NULL, NULL, napi_default, NULL},
        {"counters", NULL, Counters, NULL, NULL, NULL, napi_default, NULL},
        {"resetCounters", NULL, ResetCounters, NULL, NULL, NULL, napi_default, NULL},
        {"sync", NULL, Sync, NULL, NULL, NULL, napi_default, NULL},
        {"ticksBuffered", NULL, TicksBuffered, NULL, NULL, NULL, napi_default, NULL},
        {"profileEnable", NULL, ProfileEnable, NULL, NULL, NULL, napi_default, NULL},
        {"profileRead", NULL, ProfileRead, NULL, NULL, NULL, napi_default, NULL},
        {"setOverlap", NULL, SetOverlap, NULL, NULL, NULL, napi_default, NULL},
        {"timerStart", NULL, TimerStart, NULL, NULL, NULL, napi_default, NULL},
        {"timerStop", NULL, TimerStop, NULL, NULL, NULL, napi_default, NULL},
    };
    napi_value cls, v;
    napi_define_class(env, "ApmB200", NAPI_AUTO_LENGTH, Construct, NULL, sizeof methods / sizeof methods[0], methods, &cls);
    napi_set_named_property(env, exports, "ApmB200", cls);
    napi_create_double(env, (double)apm_abi_version(), &v);
    napi_set_named_property(env, exports, "abiVersion", v);
    const struct { const char* name; napi_callback cb; } fns[] = {
        {"quantize", Quantize}, {"ownerOfKey", OwnerOfKey}, {"ownerTableBuild", OwnerTableBuild}, {"ownerTableBuildCapped", OwnerTableBuildCapped}, {"commUniqueId", CommUniqueId},
        {"preloadKernels", PreloadKernels}, {"nameHash", NameHash}, {"calcPercentiles", CalcPercentiles}, {"partitionColumns", PartitionColumns}};
    for (size_t i = 0; i < sizeof fns / sizeof fns[0]; ++i) {
        napi_create_function(env, fns[i].name, NAPI_AUTO_LENGTH, fns[i].cb, NULL, &v);
        napi_set_named_property(env, exports, fns[i].name, v);
    }
    return exports;
}
