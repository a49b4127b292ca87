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
