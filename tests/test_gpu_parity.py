"""GPU parity: the CUDA path through the C ABI against the CPU oracle on the same seeded
streams.  Bit-exact: bucket assignment, counts, sums, min/max, percentile selection, tick
timestamps, quantised inter-stage values, signals - and, with exact-order summation, the
fp64 rolling mean and bounds too (the spec allows 1e-9 relative there)."""
import numpy as np
import pytest

from tests.gpu_util import assert_fs_equal, assert_stats_equal, run_both
from tests.streams import make_stream
from tests.test_oracle_twins import CASES

pytestmark = pytest.mark.gpu


def stream_for(c, **kw):
    return make_stream(c["seed"], c["n"], c["S"], c["J"], c["bucket_ms"], c["span"], gap_at=c["gap"],
                       shift_services=(0, 2), **kw)


@pytest.mark.parametrize("c", CASES, ids=lambda c: f"seed{c['seed']}")
def test_stream_matches_oracle(c):
    eng, orc, st, fs = run_both(c, stream_for(c))
    assert orc.ticks > 0 and st.size > 0
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())
    assert int(np.abs(fs["signal"]).sum()) > 0, "case must fire signals"
    cnt = eng.counters()
    assert cnt.ticks == orc.ticks and cnt.tx_ingested + cnt.tx_dropped == orc.ingested
    eng.close()


@pytest.mark.parametrize("c", CASES[:3], ids=lambda c: f"seed{c['seed']}")
def test_chunking_does_not_change_results(c):
    cols = stream_for(c)
    n = cols[0].size
    _, orc, st, fs = run_both(c, cols, chunks=[1, 7, n // 3, n // 3 + 1, n - 1])
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())


def test_unsorted_and_late_arrivals():
    c = dict(CASES[0], seed=11, n=3000)
    cols = make_stream(11, 3000, c["S"], c["J"], c["bucket_ms"], c["span"], late_frac=0.2, late_ms=500000)
    _, orc, st, fs = run_both(c, cols)
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())


def test_key_first_seen_through_a_too_old_transaction_is_still_emitted():
    """addData creates the (server, service) entry whatever the bucket's age
    (stream_calc_stats.js:125-126): such a key is emitted with an empty window from then on."""
    c = dict(CASES[1])
    svc, jvm, ts, el = stream_for(c)
    n = svc.size
    mid = n // 2
    svc = svc.copy(); ts = ts.copy()
    new_key = c["S"] - 1
    keep = svc != new_key                       # the key appears nowhere else ...
    svc, jvm, ts, el = svc[keep], jvm[keep], ts[keep], el[keep]
    mid = svc.size // 2
    svc = np.insert(svc, mid, new_key); jvm = np.insert(jvm, mid, 0)
    ts = np.insert(ts, mid, ts[mid] - 50 * c["bucket_ms"])  # ... except once, far older than the retained buckets
    el = np.insert(el, mid, 7)
    eng, orc, st, fs = run_both(c, (svc, jvm, ts, el))
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())
    rows = st[st["key"] == new_key]
    assert rows.size > 0 and np.all(rows["count"] == 0)
    assert eng.counters().tx_dropped >= 1


def test_empty_and_tiny_inputs():
    c = dict(CASES[0])
    z = np.zeros(0, dtype=np.int64)
    eng, orc, st, fs = run_both(c, (z, z, z, z))
    assert st.size == 0 and fs.size == 0
    one = (np.array([1], np.uint32), np.array([0], np.uint32), np.array([1578409200000], np.int64),
           np.array([5], np.int32))
    eng, orc, st, fs = run_both(c, one)
    assert st.size == 0  # the first transaction ticks with no key seen yet
    eng.close()


def test_invalid_transactions_are_counted_and_skipped():
    c = dict(CASES[1])
    svc, jvm, ts, el = stream_for(c)
    svc = svc.copy(); el = el.copy()
    svc[10] = 1000; el[20] = -5
    eng, orc, st, fs = run_both(c, (svc, jvm, ts, el))
    assert_stats_equal(st, orc.stats())
    assert eng.counters().tx_invalid == 2


def test_percentile_ties_and_boundaries():
    """n = 1, n % 4 == 0, n % 20 == 0 and heavy duplicates inside one window."""
    c = dict(seed=5, n=0, S=4, J=1, bucket_ms=10000, window=3, buffer=0, span=0, lags=(), thr=(), infl=(),
             pcts=(75.0, 95.0, 50.0, 99.0), gap=None)
    t0 = 1578409200000
    svc, ts, el = [], [], []
    for key, n in enumerate((1, 4, 20, 41)):
        for i in range(n):
            svc.append(key); ts.append(t0 + (i % 3) * 10000 + i); el.append((i * 7) % 5 + 1)
    order = np.argsort(np.array(ts), kind="stable")
    svc = np.array(svc, np.uint32)[order]; ts = np.array(ts, np.int64)[order]; el = np.array(el, np.int32)[order]
    # a closing transaction far ahead ticks with everything in the window
    svc = np.append(svc, 0).astype(np.uint32); ts = np.append(ts, t0 + 30000).astype(np.int64)
    el = np.append(el, 1).astype(np.int32)
    eng, orc, st, fs = run_both(c, (svc, np.zeros_like(svc), ts, el))
    assert_stats_equal(st, orc.stats())
    last = st[st["tick_ts"] == st["tick_ts"].max()]
    assert set(last["count"].tolist()) == {1, 4, 20, 41}


def test_acquire_commit_tick_protocol_matches_process_host():
    """The per-message protocol of consumeMsg (host finds the ticks) equals the device tick finder."""
    from apmbackend_b200 import Engine
    from tests.gpu_util import engine_config
    c = CASES[2]
    svc, jvm, ts, el = stream_for(c)
    _, orc, st, fs = run_both(c, (svc, jvm, ts, el))
    eng = Engine(engine_config(c))
    latest, start = 0, 0
    labels = ts // c["bucket_ms"]
    for i in range(svc.size):
        if labels[i] > latest:
            eng.add_batch(svc[start:i], jvm[start:i], ts[start:i], el[start:i])
            start = i
            latest = int(labels[i])
            eng.tick(latest)
    eng.add_batch(svc[start:], jvm[start:], ts[start:], el[start:])
    assert_stats_equal(eng.fetch_stats(), st)
    assert_fs_equal(eng.fetch_fs(), fs)
    eng.close()


def test_ring_load_read_and_state_roundtrip(tmp_path):
    from apmbackend_b200 import Engine
    from tests.gpu_util import engine_config
    c = CASES[0]
    cols = stream_for(c)
    n = cols[0].size
    half = n // 2
    a = Engine(engine_config(c))
    a.process_host(*[x[:half] for x in cols])
    a.release_ticks()
    path = str(tmp_path / "state.bin")
    a.state_save(path)
    b = Engine(engine_config(c))
    b.state_load(path)
    for e in (a, b):
        e.process_host(*[x[half:] for x in cols])
    assert a.fetch_stats().tobytes() == b.fetch_stats().tobytes()
    assert a.fetch_fs().tobytes() == b.fetch_fs().tobytes()
    r = a.ring_read(0, 0, 1)
    assert r.size == min(c["lags"][0], a.counters().ticks - 1) or r.size > 0
    b.ring_load(3, 1, 2, np.arange(17, dtype=np.float64))
    assert np.array_equal(b.ring_read(3, 1, 2), np.arange(17, dtype=np.float64))
    a.close(); b.close()


def test_list_loaded_for_a_silent_key_realigns_when_the_key_appears():
    """A resumed lag list of a key that stays silent for a while does not advance with the other
    keys' lists; when the key shows up, its list must continue exactly where it stopped."""
    from apmbackend_b200 import Engine
    from oracle import capi as ocapi
    from tests.gpu_util import engine_config, oracle_config
    c = dict(CASES[0], lags=(8, 40), n=3000)
    svc, jvm, ts, el = stream_for(c)
    late_key = 5
    quiet = (svc == late_key % c["S"]) & (jvm == late_key // c["S"]) & (np.arange(svc.size) < svc.size // 2)
    svc, jvm, ts, el = svc[~quiet], jvm[~quiet], ts[~quiet], el[~quiet]  # key 5 is silent in the first half
    eng = Engine(engine_config(c))
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    hist = np.round(np.linspace(20.0, 60.0, 40), 1)
    hist[7] = np.nan
    for s in range(3):
        eng.ring_load(late_key, 1, s, hist + s)
        orc.ring_load(late_key, 1, s, hist + s)
        eng.ring_load(late_key, 0, s, hist[:5] + s)   # a partial list, too
        orc.ring_load(late_key, 0, s, hist[:5] + s)
    orc.consume(svc, jvm, ts, el)
    eng.process_host(svc, jvm, ts, el)
    assert_fs_equal(eng.fetch_fs(), orc.fs())
    for li in (0, 1):
        for s in range(3):
            a, b = eng.ring_read(late_key, li, s), orc.ring_read(late_key, li, s)
            assert a.tobytes() == b.tobytes()
    eng.close()


@pytest.mark.parametrize("S,n,hi", [(50, 20000, 60000), (20000, 300000, 60000), (300, 50000, 5_000_000)],
                         ids=["few-keys", "more-keys-than-table-entries", "durations-above-2^16"])
def test_stats_only_mode_uses_the_shared_memory_tables(S, n, hi):
    """Without percentiles the keyed reduction accumulates in per-CTA shared-memory hash tables
    (misses and durations >= 2^16 ms go straight to the global planes): count, sum, min, max,
    mean, tpm and the extension stddev must still equal the oracle's."""
    from apmbackend_b200 import Engine, EngineConfig, _capi
    from oracle import capi as ocapi
    svc, jvm, ts, el = make_stream(77, n, S, 1, 60000, 12, elapsed_hi=hi, late_frac=0.02, late_ms=90000)
    if hi > 65536:
        el = el.copy()
        el[::7] = np.minimum(el[::7].astype(np.int64) * 900, hi).astype(np.int32)
    eng = Engine(EngineConfig(num_services=S, bucket_ms=60000, window=2, buffer=1, percentiles=(), lags=(), thresholds=(),
                              influences=(), flags=_capi.F_KEEP_STATS, max_batch=1 << 16, out_ticks=64))
    orc = ocapi.Pipeline(ocapi.make_config(S, 1, 60000, 2, 1, (), (0, 1), (), (), (), flags=ocapi.F_KEEP_STATS), fast=True)
    orc.consume(svc, jvm, ts, el)
    eng.process_host(svc, jvm, ts, el)
    st = eng.fetch_stats()
    assert st.size > 0 and int(st["count"].sum()) > n // 2
    assert_stats_equal(st, orc.stats())
    eng.close()


def test_true_stddev_mode_extension():
    """APM_SD_TRUE (extension, not in the reference): population stddev of the list instead of the
    reference's sqrt(mean); same operation order on both sides, so still bit for bit."""
    from apmbackend_b200 import Engine
    from oracle import capi as ocapi
    from tests.gpu_util import engine_config, oracle_config
    c = dict(CASES[0], thr=(2.5, 2.0))
    cols = stream_for(c)
    eng = Engine(engine_config(c, sd_mode=1))
    ocfg = oracle_config(c)
    ocfg.sd_mode = 1
    orc = ocapi.Pipeline(ocfg, fast=True)
    orc.consume(*cols)
    eng.process_host(*cols)
    fs = eng.fetch_fs()
    assert_fs_equal(fs, orc.fs())
    assert int(np.abs(fs["signal"]).sum()) > 0
    eng.close()


def test_bound_pinned_output_streams_the_same_fs_records():
    from apmbackend_b200 import Engine
    from tests.gpu_util import engine_config
    c = CASES[0]
    cols = stream_for(c)
    a = Engine(engine_config(c))
    a.process_host(*cols)
    want = a.fetch_fs().copy()
    a.close()
    b = Engine(engine_config(c))
    b.bind_fs_output()
    n = cols[0].size
    parts = []
    for lo in range(0, n, n // 3 + 1):
        b.process_host(*[x[lo:lo + n // 3 + 1] for x in cols])
        parts.append(b.fetch_fs(pinned=True).copy())
        b.release_ticks()
    assert np.concatenate(parts).tobytes() == want.tobytes()
    b.close()


def test_alerts_are_the_signalled_fs_records():
    c = CASES[0]
    eng, orc, st, fs = run_both(c, stream_for(c))
    eng2, _, _, _ = run_both(c, stream_for(c))
    eng.close(); eng2.close()
    from apmbackend_b200 import Engine
    from tests.gpu_util import engine_config, sort_records
    e = Engine(engine_config(c))
    e.process_host(*stream_for(c))
    al = e.fetch_alerts()
    want = fs[np.abs(fs["signal"]).sum(axis=1) > 0]
    assert al.size == want.size > 0
    assert sort_records(al, True).tobytes() == sort_records(want, True).tobytes()
    e.close()


def test_window_wider_than_a_warp():
    """51 buckets per window: the warp-per-key window kernel gives every lane two buckets."""
    c = dict(seed=21, n=20000, S=7, J=2, bucket_ms=10000, window=45, buffer=5, span=140, lags=(10, 30), thr=(2.0, 1.5),
             infl=(0.1, 0.0), pcts=(50.0, 75.0, 99.0), zpct=(1, 2), gap=(60, 66))
    _, orc, st, fs = run_both(c, stream_for(c))
    assert orc.ticks > 100
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())


def test_buckets_around_the_one_cta_sort_limit():
    """~4100 transactions per bucket: some buckets take the one-CTA bitonic sort (<= 4096 words), some the tiled
    radix sort, and the sliding window mixes runs produced by both."""
    c = dict(seed=22, n=82000, S=11, J=2, bucket_ms=10000, window=3, buffer=1, span=20, lags=(6,), thr=(1.5,),
             infl=(0.2,), pcts=(75.0, 95.0), gap=None, log_cap=1 << 14, max_batch=1 << 17)
    eng, orc, st, fs = run_both(c, stream_for(c, late_frac=0.02, late_ms=15000))
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())
    per_bucket = np.unique(stream_for(c, late_frac=0.02, late_ms=15000)[2] // 10000, return_counts=True)[1]
    assert ((per_bucket > 3000) & (per_bucket <= 4096)).any() and (per_bucket > 4096).any()
    eng.close()


def test_packed_transactions_give_the_same_records():
    """apm_process_host_packed (12 B per transaction) == apm_process_host (20 B), invalid rows included."""
    from apmbackend_b200 import Engine
    from tests.gpu_util import engine_config
    c = dict(CASES[0], n=6000, max_batch=1 << 10)   # several chunks per call
    svc, jvm, ts, el = stream_for(c, late_frac=0.05)
    svc, el = svc.copy(), el.copy()
    svc[17] = c["S"] + 3                           # unknown service: invalid in both forms
    el[29] = -5                                    # negative elapsed: invalid in both forms
    a = Engine(engine_config(c))
    ra = a.process_host(svc, jvm, ts, el)
    base = int(ts.min())
    b = Engine(engine_config(c))
    rb = b.process_host_packed((jvm.astype(np.uint32) << 16) | svc, (ts - base).astype(np.uint32), el.astype(np.int32).view(np.uint32), base)
    assert (ra.ticks, ra.tx_ingested, ra.tx_dropped, ra.tx_invalid, ra.latest_bucket) == \
           (rb.ticks, rb.tx_ingested, rb.tx_dropped, rb.tx_invalid, rb.latest_bucket) and ra.tx_invalid == 2
    assert a.fetch_stats().tobytes() == b.fetch_stats().tobytes() and a.fetch_fs().tobytes() == b.fetch_fs().tobytes()
    a.close()
    b.close()


def test_full_output_ring_refuses_the_chunk_and_leaves_the_state_untouched():
    """APM_E_WOULDBLOCK (the `pause` path of stream_calc_stats.js:270-276): a chunk with more ticks than the
    output ring has room for is refused before anything of it is ingested; fed in smaller pieces with a
    fetch in between, the same stream gives the oracle's records."""
    from apmbackend_b200 import Engine
    from apmbackend_b200._capi import ApmError
    from oracle import capi as ocapi
    from tests.gpu_util import engine_config, oracle_config
    c = dict(CASES[2], out_ticks=6)
    svc, jvm, ts, el = stream_for(c)
    eng = Engine(engine_config(c))
    with pytest.raises(ApmError) as e:
        eng.process_host(svc, jvm, ts, el)            # ~40 ticks, room for 6
    assert e.value.status == 4 and eng.counters().ticks == 0 and eng.counters().tx_ingested == 0
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    orc.consume(svc, jvm, ts, el)
    st_parts, fs_parts = [], []
    step = svc.size // 40
    for a in range(0, svc.size, step):
        b = min(svc.size, a + step)
        eng.process_host(svc[a:b], jvm[a:b], ts[a:b], el[a:b])
        if eng.ticks_buffered >= 3:
            st_parts.append(eng.fetch_stats().copy()); fs_parts.append(eng.fetch_fs().copy())
            eng.release_ticks()
    st_parts.append(eng.fetch_stats().copy()); fs_parts.append(eng.fetch_fs().copy())
    assert_stats_equal(np.concatenate(st_parts), orc.stats())
    assert_fs_equal(np.concatenate(fs_parts), orc.fs())
    eng.close()


def test_refused_chunk_in_the_middle_of_a_long_call_reports_what_was_consumed():
    from apmbackend_b200 import Engine
    from apmbackend_b200._capi import ApmError
    from oracle import capi as ocapi
    from tests.gpu_util import engine_config, oracle_config
    c = dict(CASES[2], out_ticks=12, max_batch=256)     # chunks of 16 .. 64 transactions, ~40 ticks in the stream
    svc, jvm, ts, el = stream_for(c)
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    orc.consume(svc, jvm, ts, el)
    eng = Engine(engine_config(c))
    st_parts, fs_parts, off, refusals = [], [], 0, 0
    while off < svc.size:
        try:
            eng.process_host(svc[off:], jvm[off:], ts[off:], el[off:])
            off = svc.size
        except ApmError as e:
            assert e.status == 4
            r = e.result
            used = r.tx_ingested + r.tx_dropped + r.tx_invalid
            assert used > 0 or eng.ticks_buffered > 0
            off += used
            refusals += 1
        st_parts.append(eng.fetch_stats().copy()); fs_parts.append(eng.fetch_fs().copy())
        eng.release_ticks()
    assert refusals >= 2
    assert_stats_equal(np.concatenate(st_parts), orc.stats())
    assert_fs_equal(np.concatenate(fs_parts), orc.fs())
    eng.close()


def test_capacity_errors_name_the_limit():
    """A bucket with more transactions than max_tx_per_bucket and a batch above max_batch are APM_E_CAPACITY with the limit's name in the message - never silent truncation."""
    from apmbackend_b200 import Engine
    from apmbackend_b200._capi import ApmError
    from tests.gpu_util import engine_config
    c = dict(CASES[2])
    svc, jvm, ts, el = stream_for(c)
    eng = Engine(engine_config(dict(c, log_cap=16)))            # ~60 transactions per bucket, room for 16
    with pytest.raises(ApmError, match="max_tx_per_bucket") as e:
        eng.process_host(svc, jvm, ts, el)
    assert e.value.status == 3
    eng.close()
    eng = Engine(engine_config(dict(c, max_batch=64)))
    import ctypes
    from apmbackend_b200._capi import TxColumns
    pin = TxColumns()
    assert eng._lib.apm_ingest_acquire(eng._h, 64, ctypes.byref(pin)) == 0
    with pytest.raises(ApmError, match="max_batch") as e:
        eng._check(eng._lib.apm_ingest_acquire(eng._h, 65, ctypes.byref(pin)))
    assert e.value.status == 3
    eng.close()
