"""The narrow host forms (12 and 8 bytes per transaction) against the 20-byte columns: same ticks, same records,
same counters - on small streams with every edge the forms have (invalid rows, escapes, deltas at the top of their
field, bases far from the first bucket edge) and once at BASELINE size (100 M transactions, the multi-chunk ramp the
end-to-end bench leg goes through)."""
import numpy as np
import pytest

from apmbackend_b200 import Engine, _capi
from apmbackend_b200 import synthetic as syn
from apmbackend_b200.engine import pack8
from tests.gpu_util import engine_config
from tests.streams import make_stream

pytestmark = pytest.mark.gpu


def _results(eng):
    out = (eng.fetch_stats().copy(), eng.fetch_fs().copy(), eng.fetch_alerts().copy())
    c = eng.counters()
    eng.release_ticks()
    return out, (c.tx_ingested, c.tx_dropped, c.tx_invalid, c.ticks)


def _same(a, b):
    for x, y in zip(a[0], b[0]):
        assert x.tobytes() == y.tobytes()
    assert a[1] == b[1]


CFG = dict(S=50, J=3, bucket_ms=10000, window=3, buffer=1, lags=(6, 15), thr=(1.5, 2.5), infl=(0.3, 0.0), pcts=(75.0, 95.0),
           max_batch=1 << 10)


def _edge_stream():
    svc, jvm, ts, el = make_stream(81, 7000, CFG["S"], CFG["J"], CFG["bucket_ms"], 100, late_frac=0.05, shift_services=(0, 1, 4))
    svc, jvm, el = svc.copy(), jvm.copy(), el.copy()
    svc[[10, 3000]] = CFG["S"] + 7           # ids out of range
    jvm[[20, 6999]] = CFG["J"]
    el[[30, 4000]] = -5                      # negative durations: invalid
    el[[40, 41, 5000]] = [0xFFFFFF, 0xFFFFFF + 1, 2_000_000_000]   # durations at and above the 24-bit field
    el[50] = 0xFFFFFE
    return svc, jvm, ts, el


@pytest.mark.parametrize("slack", [0, 1234, "top"])
def test_packed8_equals_the_wide_columns(slack):
    svc, jvm, ts, el = _edge_stream()
    span = int(ts.max() - ts.min())
    base = int(ts.min()) - (((1 << 24) - 1 - span) if slack == "top" else slack)  # "top": the last delta is 2^24 - 1
    wide = Engine(engine_config(CFG))
    wide.process_host(svc, jvm, ts, el)
    want = _results(wide)
    assert want[1][2] >= 6  # the invalid rows were counted
    words, esc = pack8(svc, jvm, ts, el, CFG["S"], CFG["J"], base)
    assert esc.size == 5 and (words >> np.uint64(48)).max() == 0xFFFF
    if slack == "top":
        assert int((words >> np.uint64(24)).max() & 0xFFFFFF) == (1 << 24) - 1
    nar = Engine(engine_config(CFG))
    n = svc.size
    cuts = [0, 1, n // 3, n]  # several calls: escape indices are relative to the call
    for a, b in zip(cuts[:-1], cuts[1:]):
        e = esc[(esc["index"] >= a) & (esc["index"] < b)].copy()
        e["index"] -= a
        nar.process_host_packed8(words[a:b], e, base)
    _same(_results(nar), want)
    wide.close(); nar.close()


def test_packed8_all_ones_field_without_an_escape_is_a_duration():
    c = dict(CFG, S=4, J=1)
    ts = (1578409200000 + np.arange(0, 60000, 100)).astype(np.int64)
    n = ts.size
    svc = (np.arange(n) % 4).astype(np.uint32)
    jvm = np.zeros(n, np.uint32)
    el = np.full(n, 100, np.int32)
    el[7] = 0xFFFFFF
    a = Engine(engine_config(c)); b = Engine(engine_config(c))
    a.process_host(svc, jvm, ts, el)
    words, _ = pack8(svc, jvm, ts, el, 4, 1, int(ts[0]))
    b.process_host_packed8(words, None, int(ts[0]))  # the list is dropped: the field stands for itself
    _same(_results(b), _results(a))
    a.close(); b.close()


def test_packed8_rejects_bad_arguments():
    eng = Engine(engine_config(dict(CFG, S=4, J=1)))
    w = np.zeros(4, np.uint64)
    esc = np.zeros(2, _capi.ESCAPE_DTYPE)
    esc["index"] = [3, 1]
    with pytest.raises(_capi.ApmError):
        eng.process_host_packed8(w, esc, 1578409200000)   # not ascending
    with pytest.raises(_capi.ApmError):
        eng.process_host_packed8(w, None, -1)
    eng.close()


def test_narrow_forms_at_baseline_size_100m_transactions():
    """100 M transactions, 10 k keys, 60 one-minute buckets through apm_process_host (20 B), apm_process_host_packed
    (12 B, base 2^32 - 1 - span before the first timestamp) and apm_process_host_packed8 (8 B, base 2^24 - 1 - span
    before it), each in the ramped multi-chunk form the bench's end-to-end leg uses (max_batch 2^25): byte-equal
    records and counters."""
    n, K, span_buckets, bucket_ms = 100_000_000, 10_000, 60, 60_000
    spec = syn.StreamSpec(seed=21, n=n, num_services=K, num_jvms=1, span_ms=span_buckets * bucket_ms, shift_mod=97)
    stream = syn.SynthStream(spec, device=True)
    dc = syn.DeviceColumns(n)
    stream.generate_device(dc)
    svc, jvm, ts, el = dc.to_host()
    dc.free(); stream.free()
    rng = np.random.default_rng(5)
    bad = rng.integers(0, n, 2000)
    el[bad[:500]] = -1
    svc[bad[500:1000]] = K + 3
    el[bad[1000:1500]] = 0xFFFFFF + rng.integers(0, 1000, 500).astype(np.int32)   # above the 24-bit field, 25 bits in the log
    el[bad[1500:]] = 0xFFFFFF
    c = dict(S=K, J=1, bucket_ms=bucket_ms, window=1, buffer=0, lags=(8,), thr=(3.0,), infl=(0.1,), pcts=(50.0, 95.0, 99.0),
             zpct=(1, 2), max_batch=1 << 25, log_cap=int(n / span_buckets * 1.5), out_ticks=span_buckets + 4, max_alerts=3 * K)
    results = []
    span = int(ts.max() - ts.min())
    for form in ("wide", "packed12", "packed8"):
        eng = Engine(engine_config(c))
        if form == "wide":
            eng.process_host(svc, jvm, ts, el)
        elif form == "packed12":
            base = int(ts.min()) - ((1 << 32) - 1 - span)
            ok = (svc < K)
            key = np.where(ok, svc, 0xFFFF).astype(np.uint32)  # jvm 0; an id >= num_services stays out of range
            eng.process_host_packed(key, (ts - base).astype(np.uint32), el.view(np.uint32), base)
        else:
            base = int(ts.min()) - ((1 << 24) - 1 - span)
            words, esc = pack8(svc, jvm, ts, el, K, 1, base)
            assert esc.size >= 1000
            eng.process_host_packed8(words, esc, base)
        results.append(_results(eng))
        eng.close()
    assert results[0][1][3] >= span_buckets - 1
    assert results[0][1][2] >= 1000
    _same(results[1], results[0])
    _same(results[2], results[0])
