"""The Node-API shim (napi/apm_b200_addon.c) against a fake Node-API: module registration and
argument checking run anywhere; the data path runs on the GPU box and must hand JavaScript the
same record bytes as the ctypes path."""
import numpy as np
import pytest

from apmbackend_b200 import _capi
from tests.napi_driver import FakeNode, JsError

OPTS = dict(numServices=6, numJvms=2, bucketMs=10000, window=30, buffer=6, percentiles=[75, 95], lags=[8, 40],
            thresholds=[2.0, 1.5], influences=[0.1, 0.0], outTicks=256, maxBatch=1 << 16, maxTxPerBucket=1 << 16)


def test_module_registers_the_class_and_its_methods():
    node = FakeNode()
    assert node.L.fk_number_value(node.prop(node.exports, "abiVersion")) == 1.0
    cls = node.prop(node.exports, "ApmB200")
    for m in ("addBatch", "tick", "processColumns", "processPacked", "setKeyParams", "saveState", "loadState", "dictSet", "parseLog",
              "feedParsed", "close"):
        assert node.prop(cls, m), m


def test_constructor_errors_become_js_exceptions():
    node = FakeNode()
    with pytest.raises(JsError, match="num_services"):
        node.new("ApmB200", node.obj(dict(OPTS, numServices=0)))
    with pytest.raises(JsError, match="at most 4"):
        node.new("ApmB200", node.obj(dict(OPTS, percentiles=[1, 2, 3, 4, 5])))
    if _capi.device_count() == 0:
        with pytest.raises(JsError, match="no CPU fallback"):
            node.new("ApmB200", node.obj(OPTS))


@pytest.mark.gpu
def test_records_through_the_shim_equal_the_ctypes_path(tmp_path):
    from apmbackend_b200 import Engine, EngineConfig
    from tests.streams import make_stream
    svc, jvm, ts, el = make_stream(41, 3000, 6, 2, 10000, 90, shift_services=(0, 2))
    eng = Engine(EngineConfig(num_services=6, num_jvms=2, lags=(8, 40), thresholds=(2.0, 1.5), influences=(0.1, 0.0),
                              out_ticks=256, max_batch=1 << 16, max_tx_per_bucket=1 << 16))
    eng.process_host(svc, jvm, ts, el)
    want_st, want_fs, want_al = eng.fetch_stats().copy(), eng.fetch_fs().copy(), eng.fetch_alerts().copy()
    eng.close()

    node = FakeNode()
    gpu = node.new("ApmB200", node.obj(OPTS))
    res = node.call(gpu, "processColumns", node.typed(svc), node.typed(jvm), node.typed(ts), node.typed(el))
    assert node.L.fk_number_value(node.prop(res, "ticks")) > 50
    assert node.buffer(node.prop(res, "stats"), _capi.STAT_DTYPE).tobytes() == want_st.tobytes()
    assert node.buffer(node.prop(res, "fs"), _capi.FS_DTYPE).tobytes() == want_fs.tobytes()
    assert node.buffer(node.prop(res, "alerts"), _capi.FS_DTYPE).tobytes() == want_al.tobytes()

    # the 12-byte form of the same transactions
    gpu3 = node.new("ApmB200", node.obj(OPTS))
    base = int(ts.min())
    res3 = node.call(gpu3, "processPacked", node.typed((jvm.astype(np.uint32) << 16) | svc), node.typed((ts - base).astype(np.uint32)),
                     node.typed(el.view(np.uint32)), node.number(base))
    assert node.buffer(node.prop(res3, "stats"), _capi.STAT_DTYPE).tobytes() == want_st.tobytes()
    assert node.buffer(node.prop(res3, "fs"), _capi.FS_DTYPE).tobytes() == want_fs.tobytes()
    node.call(gpu3, "close")

    # the per-message protocol: addBatch up to the tick, tick(label), as consumeMsg does (stream_calc_stats.js:348-370)
    gpu2 = node.new("ApmB200", node.obj(OPTS))
    labels = ts // 10000
    latest, start, st_parts = 0, 0, []
    for i in range(svc.size):
        if labels[i] > latest:
            if i > start:
                node.call(gpu2, "addBatch", node.typed(svc[start:i]), node.typed(jvm[start:i]), node.typed(ts[start:i]),
                          node.typed(el[start:i]))
            start, latest = i, int(labels[i])
            r = node.call(gpu2, "tick", node.number(latest))
            st_parts.append(node.buffer(node.prop(r, "stats"), _capi.STAT_DTYPE))
    assert np.concatenate(st_parts).tobytes() == want_st.tobytes()

    # argument checking and state files
    with pytest.raises(JsError, match="equal length"):
        node.call(gpu2, "addBatch", node.typed(svc[:5]), node.typed(jvm[:4]), node.typed(ts[:5]), node.typed(el[:5]))
    with pytest.raises(JsError, match="must exceed"):
        node.call(gpu2, "tick", node.number(1))
    path = str(tmp_path / "state.bin")
    node.call(gpu2, "saveState", node.string(path))
    node.call(gpu2, "loadState", node.string(path))
    node.call(gpu2, "setKeyParams", node.number(3), node.number(0), node.number(25.0), node.number(0.1))
    node.call(gpu2, "close")
    with pytest.raises(JsError, match="closed"):
        node.call(gpu2, "tick", node.number(latest + 1))


@pytest.mark.gpu
def test_log_feeder_through_the_shim_equals_the_ctypes_path():
    from apmbackend_b200 import Engine, EngineConfig
    from apmbackend_b200.transaction_parser import rewrite_service
    from tests.logs import SERVICES, make_log
    text = make_log(seed=31, n=1500, span_ms=400_000, unique_ids=True)
    names = {}
    for sv in SERVICES + ["Svc", "undefined", "Login"]:
        for nm in (rewrite_service(sv), rewrite_service("S:" + sv)):
            names.setdefault(nm, len(names))
    cfg = dict(num_services=64, num_jvms=2, lags=(8, 40), thresholds=(2.0, 1.5), influences=(0.1, 0.0), out_ticks=256,
               max_batch=1 << 16, max_tx_per_bucket=1 << 16)
    eng = Engine(EngineConfig(**cfg))
    eng.dict_set(names)
    want_rec, want_res = eng.parse_log(text, 1, 1)
    eng.feed_parsed()
    want_st, want_fs = eng.fetch_stats().copy(), eng.fetch_fs().copy()
    eng.close()

    node = FakeNode()
    gpu = node.new("ApmB200", node.obj(dict(OPTS, numServices=64)))
    node.call(gpu, "dictSet", node.strings(list(names)), node.typed(np.array(list(names.values()), np.uint32)))
    r = node.call(gpu, "parseLog", node.typed(np.frombuffer(text, np.uint8)), node.number(1), node.number(1),
                  node.numbers([-9.3e18, 0]))
    assert node.buffer(node.prop(r, "records"), _capi.LOG_DTYPE).tobytes() == want_rec.tobytes() and len(want_rec) > 1000
    assert node.L.fk_number_value(node.prop(r, "fed")) == want_res.fed > 500
    assert node.L.fk_number_value(node.prop(r, "unknownService")) == 0
    out = node.call(gpu, "feedParsed")
    assert node.buffer(node.prop(out, "stats"), _capi.STAT_DTYPE).tobytes() == want_st.tobytes()
    assert node.buffer(node.prop(out, "fs"), _capi.FS_DTYPE).tobytes() == want_fs.tobytes()
    with pytest.raises(JsError, match="equal length"):
        node.call(gpu, "dictSet", node.strings(["a", "b"]), node.typed(np.array([1], np.uint32)))
    with pytest.raises(JsError, match="Uint8Array"):
        node.call(gpu, "parseLog", node.typed(np.zeros(4, np.uint32)), node.number(0), node.number(0))
    node.call(gpu, "close")


# ---- every entry point of include/apm_b200.h has a JavaScript face ---------------------------------------------------
# the C ABI's symbols and the addon method / function that reaches each (None: takes a device pointer, which JavaScript
# does not have; the host-array form beside it is the one the addon uses)
ABI_TO_JS = {
    "apm_abi_version": "abiVersion", "apm_ctx_create": "new ApmB200", "apm_ctx_destroy": "close", "apm_last_error": "(exception text)",
    "apm_ingest_acquire": "addBatch", "apm_ingest_commit": "addBatch", "apm_tick": "tick", "apm_process_host": "processColumns",
    "apm_process_device": None, "apm_process_host_packed": "processPacked", "apm_process_host_packed8": "processPacked8",
    "apm_set_tick_marks": "setTickMarks", "apm_set_key_params": "setKeyParams", "apm_set_alert_rules": "setAlertRules",
    "apm_set_alert_key_params": "setAlertKeyParams", "apm_drop_lag": "dropLag", "apm_fetch_stats": "tick", "apm_fetch_fs": "tick",
    "apm_fetch_alerts": "tick", "apm_release_ticks": "tick", "apm_fetch_passthrough": "tick", "apm_host_alloc": "bindFsOutput",
    "apm_host_free": "close", "apm_bind_fs_output": "bindFsOutput", "apm_ticks_buffered": "ticksBuffered", "apm_sync": "sync",
    "apm_ring_load": "ringLoad", "apm_ring_read": "ringRead", "apm_ring_load_all": "ringLoadAll", "apm_state_save": "saveState",
    "apm_state_load": "loadState", "apm_profile_enable": "profileEnable", "apm_profile_read": "profileRead", "apm_set_overlap": "setOverlap",
    "apm_timer_start": "timerStart", "apm_timer_stop": "timerStop", "apm_set_latest": "setLatest", "apm_mark_seen": "markSeen",
    "apm_log_read": "logRead", "apm_latest": "latest", "apm_get_counters": "counters", "apm_reset_counters": "resetCounters",
    "apm_preload_kernels": "preloadKernels", "apm_quantize": "quantize", "apm_sorted_percentiles": None,
    "apm_sorted_percentiles_host": "calcPercentiles", "apm_sorted_percentiles_noalloc": None, "apm_owner_of_key": "ownerOfKey",
    "apm_owner_table_build": "ownerTableBuild", "apm_owner_table_build_capped": "ownerTableBuildCapped", "apm_partition_host": "partitionColumns", "apm_comm_unique_id": "commUniqueId",
    "apm_comm_init": "commInit", "apm_comm_set_capacity": "commSetCapacity", "apm_bind_gather_output": "bindGatherOutput",
    "apm_gather_alerts": "gatherAlerts", "apm_name_hash": "nameHash", "apm_dict_set": "dictSet", "apm_parse_log_host": "parseLog",
    "apm_feed_parsed": "feedParsed",
}
MODULE_FUNCTIONS = ("quantize", "ownerOfKey", "ownerTableBuild", "ownerTableBuildCapped", "commUniqueId", "preloadKernels", "nameHash", "calcPercentiles",
                    "partitionColumns")


def test_every_abi_entry_point_is_reachable_from_javascript():
    assert set(ABI_TO_JS) == set(_capi.SYMBOLS), set(ABI_TO_JS) ^ set(_capi.SYMBOLS)
    node = FakeNode()
    cls = node.prop(node.exports, "ApmB200")
    src = open(__file__.replace("tests/test_napi_shim.py", "napi/apm_b200_addon.c")).read()
    for sym, js in ABI_TO_JS.items():
        if js is None:
            continue
        assert sym + "(" in src, f"the addon never calls {sym}"
        if js.startswith(("new ", "(")) or js == "abiVersion":
            continue
        assert node.prop(node.exports if js in MODULE_FUNCTIONS else cls, js), js


def test_module_level_helpers_equal_the_c_abi():
    from apmbackend_b200 import sharding
    from tests.streams import make_stream
    lib = _capi.lib()
    node = FakeNode()
    for x, d in ((33.33333333333333, 1), (0.05, 1), (1.45, 1), (8.345, 2), (1234.05, 1)):
        assert node.L.fk_number_value(node.fn("quantize", node.number(x), node.number(d))) == lib.apm_quantize(x, d)
    for k in (0, 1, 12345, 79999):
        assert node.L.fk_number_value(node.fn("ownerOfKey", node.number(k), node.number(8))) == lib.apm_owner_of_key(k, 8)
    counts = np.r_[np.full(10, 80_000_000), np.full(990, 20_020)].astype(np.uint64)
    got = node.buffer(node.fn("ownerTableBuild", node.typed(counts), node.number(8), node.number(160_000)), np.uint32)
    assert np.array_equal(got, sharding.balanced_owner_table(counts, 8, key_cost=160_000))
    got = node.buffer(node.fn("ownerTableBuildCapped", node.typed(counts), node.number(8), node.number(0.01)), np.uint32)
    assert np.array_equal(got, sharding.capped_owner_table(counts, 8, 0.01))
    assert node.string_value(node.fn("nameHash", node.string("S:getFoo"))) == str(lib.apm_name_hash(b"S:getFoo", 8))
    # host binning: the same sub-streams and marks as the ctypes path
    svc, jvm, ts, el = make_stream(91, 3000, 12, 2, 10000, 60, late_frac=0.05)
    want, latest = sharding.partition_native(svc, jvm, ts, el, 10000, 12, 2, 3)
    opts = node.obj(dict(bucketMs=10000, latest=0, numServices=12, numJvms=2, nranks=3))
    res = node.fn("partitionColumns", node.typed(svc), node.typed(jvm), node.typed(ts), node.typed(el), opts)
    assert node.L.fk_number_value(node.prop(res, "latest")) == latest
    for r in range(3):
        part = node.prop(res, str(r))
        for name, dt, col in (("svc", np.uint32, 0), ("jvm", np.uint32, 1), ("endTs", np.int64, 2), ("elapsed", np.int32, 3)):
            assert np.array_equal(node.buffer(node.prop(part, name), dt), want[r][0][col])
        assert node.buffer(node.prop(part, "marks"), _capi.MARK_DTYPE).tobytes() == want[r][1].tobytes()
    with pytest.raises(JsError, match="partitionColumns"):
        node.fn("partitionColumns", node.typed(svc), node.typed(jvm), node.typed(ts), node.typed(el), node.obj(dict(bucketMs=10000)))


@pytest.mark.gpu
def test_the_rest_of_the_abi_through_the_shim_equals_the_ctypes_path():
    from apmbackend_b200 import Engine, EngineConfig
    from apmbackend_b200.engine import pack8
    from tests.streams import make_stream
    S, J = 6, 2
    svc, jvm, ts, el = make_stream(43, 4000, S, J, 10000, 120, shift_services=(0, 2), late_frac=0.05)
    cfg = dict(num_services=S, num_jvms=J, lags=(8, 40), thresholds=(2.0, 1.5), influences=(0.1, 0.0), out_ticks=256,
               max_batch=1 << 16, max_tx_per_bucket=1 << 16)
    pt_flags = (_capi.F_PERCENTILES | _capi.F_ZSCORE | _capi.F_KEEP_STATS | _capi.F_KEEP_FS | _capi.F_ALERTS | _capi.F_PASSTHROUGH)
    half = svc.size // 2
    lists = np.round(np.random.default_rng(3).uniform(50, 400, (3, 8)), 1)

    # ---- ctypes path: lists loaded, alert rules on, per-key settings, first half as 8-byte words, second half with marks
    eng = Engine(EngineConfig(flags=pt_flags, **cfg))
    for s_ in range(3):
        eng.ring_load(5, 0, s_, lists[s_])
    eng.set_alert_rules(window=6, required=2, alert_on_both_only=False, hard_max_ms=900.0, hard_min_ms=10.0, hard_min_tpm=0.1)
    eng.set_alert_key_params(0, np.r_[300.0, np.zeros(S * J - 1)], np.r_[np.zeros(S * J - 1), 1].astype(np.uint8))
    eng.set_key_params(3, 0, [25.0], [0.1])
    base = int(ts.min())
    words, esc = pack8(svc[:half], jvm[:half], ts[:half], el[:half], S, J, base)
    eng.process_host_packed8(words, esc, base)
    want1 = (eng.fetch_stats().copy(), eng.fetch_fs().copy(), eng.fetch_alerts().copy(), eng.fetch_passthrough())
    eng.release_ticks()
    labels = ts[half:] // 10000
    run = np.maximum.accumulate(np.r_[eng.latest, labels])
    pos = np.flatnonzero(run[1:] > run[:-1])
    marks = np.zeros(pos.size, _capi.MARK_DTYPE)
    marks["pos"], marks["label"] = pos, labels[pos]
    eng.set_tick_marks(marks)
    eng.process_host(svc[half:], jvm[half:], ts[half:], el[half:])
    want2 = (eng.fetch_stats().copy(), eng.fetch_fs().copy(), eng.fetch_alerts().copy(), eng.fetch_passthrough())
    eng.release_ticks()
    want_ring = eng.ring_read(5, 0, 1)
    want_log = eng.log_read(eng.latest)
    want_cnt = eng.counters()
    eng.drop_lag(1)
    assert eng.ring_read(5, 1, 0).size == 0
    eng.close()

    # ---- the same through the addon
    node = FakeNode()
    gpu = node.new("ApmB200", node.obj(dict(OPTS, flags=pt_flags)))
    node.fn("preloadKernels", node.number(0))
    node.call(gpu, "bindFsOutput")
    for s_ in range(3):
        node.call(gpu, "ringLoad", node.number(5), node.number(0), node.number(s_), node.typed(lists[s_].copy()))
    node.call(gpu, "setAlertRules", node.obj(dict(window=6, required=2, alertOnBothOnly=0, hardMaxMs=900.0, hardMinMs=10.0, hardMinTpm=0.1)))
    node.call(gpu, "setAlertKeyParams", node.number(0), node.typed(np.r_[300.0, np.zeros(S * J - 1)]),
              node.typed(np.r_[np.zeros(S * J - 1), 1].astype(np.uint8)))
    node.call(gpu, "setKeyParamsRange", node.number(3), node.number(0), node.typed(np.array([25.0])), node.typed(np.array([0.1])))
    esc_flat = np.stack([esc["index"].astype(np.int32), esc["elapsed_ms"]], axis=1).reshape(-1).astype(np.int32)
    node.call(gpu, "profileEnable", node.number(1))
    node.call(gpu, "timerStart")
    r1 = node.call(gpu, "processPacked8", node.typed(words), node.typed(esc_flat), node.number(base))
    assert node.L.fk_number_value(node.call(gpu, "timerStop")) > 0
    prof = node.call(gpu, "profileRead")
    assert node.buffer(node.prop(prof, "ms"), np.float64).sum() > 0
    node.call(gpu, "setOverlap", node.number(1))
    assert node.L.fk_number_value(node.call(gpu, "latest")) == int(ts[:half].max() // 10000)
    node.call(gpu, "setTickMarks", node.typed(np.stack([marks["pos"].astype(np.int64), marks["label"]], axis=1).reshape(-1)))
    r2 = node.call(gpu, "processColumns", node.typed(svc[half:]), node.typed(jvm[half:]), node.typed(ts[half:]), node.typed(el[half:]))
    for res, want in ((r1, want1), (r2, want2)):
        assert node.buffer(node.prop(res, "stats"), _capi.STAT_DTYPE).tobytes() == want[0].tobytes()
        assert node.buffer(node.prop(res, "fs"), _capi.FS_DTYPE).tobytes() == want[1].tobytes()
        assert node.buffer(node.prop(res, "alerts"), _capi.FS_DTYPE).tobytes() == want[2].tobytes()
        pt = node.prop(res, "passthrough")
        assert np.array_equal(node.buffer(node.prop(pt, "endTs"), np.int64), want[3][0])
        assert np.array_equal(node.buffer(node.prop(pt, "seq"), np.uint32), want[3][1])
        assert np.array_equal(node.buffer(node.prop(pt, "tickCounts"), np.uint32), want[3][2])
    assert want1[2]["flags"].max() >= 3, "the alert rules must have triggered"
    assert np.array_equal(node.buffer(node.call(gpu, "ringRead", node.number(5), node.number(0), node.number(1)), np.float64), want_ring)
    assert np.array_equal(node.buffer(node.call(gpu, "logRead", node.number(int(ts.max() // 10000))), np.uint64), want_log)
    cnt = node.call(gpu, "counters")
    for js, f in (("txIngested", "tx_ingested"), ("txDropped", "tx_dropped"), ("ticks", "ticks"), ("keysSeen", "keys_seen"),
                  ("fsRecords", "fs_records"), ("alerts", "alerts")):
        assert node.L.fk_number_value(node.prop(cnt, js)) == getattr(want_cnt, f), js
    node.call(gpu, "dropLag", node.number(1))
    assert node.buffer(node.call(gpu, "ringRead", node.number(5), node.number(1), node.number(0)), np.float64).size == 0
    node.call(gpu, "sync")
    assert node.L.fk_number_value(node.call(gpu, "ticksBuffered")) == 0
    node.call(gpu, "resetCounters")
    assert node.L.fk_number_value(node.prop(node.call(gpu, "counters"), "ticks")) == 0
    node.call(gpu, "close")

    # resume-file interop: setLatest / markSeen / ringLoadAll on a fresh context
    gpu2 = node.new("ApmB200", node.obj(OPTS))
    node.call(gpu2, "setLatest", node.number(157840920))
    node.call(gpu2, "markSeen", node.number(2), node.number(3))
    allv = np.round(np.random.default_rng(4).uniform(10, 90, (8, 3, S * J)), 1)
    node.call(gpu2, "ringLoadAll", node.number(0), node.typed(allv.reshape(-1).copy()), node.number(8))
    assert np.array_equal(node.buffer(node.call(gpu2, "ringRead", node.number(7), node.number(0), node.number(2)), np.float64), allv[:, 2, 7])
    r = node.call(gpu2, "tick", node.number(157840921))
    st = node.buffer(node.prop(r, "stats"), _capi.STAT_DTYPE)
    assert sorted(st["key"].tolist()) == [2, 3, 4]
    with pytest.raises(JsError, match="ringLoadAll"):
        node.call(gpu2, "ringLoadAll", node.number(0), node.typed(allv.reshape(-1)[:10].copy()), node.number(8))
    node.call(gpu2, "close")

    # Array.prototype.binaryConcat + calcPercentile for one host array
    vals = np.random.default_rng(5).integers(0, 5000, 12345).astype(np.int32)
    res = node.fn("calcPercentiles", node.typed(vals), node.numbers([50, 75, 95]))
    assert np.array_equal(node.buffer(node.prop(res, "sorted"), np.int32), np.sort(vals))
    from oracle import capi as ocapi
    want_p = [ocapi.calc_percentile(np.sort(vals).tolist(), p) for p in (50, 75, 95)]
    assert node.buffer(node.prop(res, "pct"), np.float64).tolist() == want_p

    # the gather on a world of one rank: its own alerts, through NCCL - on demand, then streamed tick by tick
    eng = Engine(EngineConfig(**cfg))
    eng.process_host(svc, jvm, ts, el)
    want_al = eng.fetch_alerts().copy()
    eng.close()
    assert want_al.size > 0
    for bound in (False, True):
        g3 = node.new("ApmB200", node.obj(OPTS))
        uid = node.buffer(node.fn("commUniqueId"), np.uint8)
        assert uid.size == 128
        node.call(g3, "commInit", node.number(0), node.number(1), node.typed(uid))
        node.call(g3, "commSetCapacity", node.number(1 << 14))
        if bound:
            node.call(g3, "bindGatherOutput", node.number(1 << 14))
        res = node.call(g3, "processColumns", node.typed(svc), node.typed(jvm), node.typed(ts), node.typed(el))
        assert node.buffer(node.prop(res, "alerts"), _capi.FS_DTYPE).tobytes() == want_al.tobytes()
        assert node.buffer(node.prop(res, "gathered"), _capi.FS_DTYPE).tobytes() == want_al.tobytes()  # rank 0: pad_[2] stays 0
        assert node.buffer(node.call(g3, "gatherAlerts", node.number(1 << 14)), _capi.FS_DTYPE).size == 0  # nothing retained any more
        node.call(g3, "close")
