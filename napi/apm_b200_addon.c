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

/* { ticks, latestBucket, stats, fs, alerts } with the records of the retained ticks, then releases them */
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
        apm_fs_record* tmp = (apm_fs_record*)malloc((cap ? cap : 1) * sizeof *tmp);
        apm_status st = apm_fetch_fs(o->ctx, tmp, cap, &n);
        if (st) { free(tmp); return fail(env, o->ctx, st, "apm_fetch_fs"); }
        napi_create_arraybuffer(env, n * sizeof *tmp, &p, &v);
        memcpy(p, tmp, n * sizeof *tmp);
        free(tmp);
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

static napi_value Close(napi_env env, napi_callback_info info) {
    size_t argc = 0;
    napi_value self;
    addon_obj* o = NULL;
    napi_get_cb_info(env, info, &argc, NULL, &self, NULL);
    if (napi_unwrap(env, self, (void**)&o) == napi_ok && o && o->ctx) {
        apm_ctx_destroy(o->ctx);
        o->ctx = NULL;
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
    };
    napi_value cls, v;
    napi_define_class(env, "ApmB200", NAPI_AUTO_LENGTH, Construct, NULL, sizeof methods / sizeof methods[0], methods, &cls);
    napi_set_named_property(env, exports, "ApmB200", cls);
    napi_create_double(env, (double)apm_abi_version(), &v);
    napi_set_named_property(env, exports, "abiVersion", v);
    return exports;
}
