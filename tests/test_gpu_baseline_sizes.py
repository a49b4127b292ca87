"""Parity at BASELINE.json's sizes.

cfg1 (configs[0]): 1 M transactions, 200 services x 2 JVMs, the reference's own settings
(10 s buckets, window 30, buffer 6, p75/p95, lags 360 and 8640) - every record of every tick
against the oracle, bit for bit.

cfg3 (configs[2]): 100 M transactions, 10 k keys, 1-min buckets, p50/p95/p99 + z-score against
a pre-filled 10 080-slot baseline - the oracle follows a 1 % sample of the keys through the
whole stream (ticks stay global), and size-independent invariants cover the rest: every
transaction is accounted for, and per tick the counts over all keys equal the number of
transactions the closing bucket held when the tick fired."""
import numpy as np
import pytest

from apmbackend_b200 import Engine, EngineConfig, _capi
from apmbackend_b200 import synthetic as syn
from oracle import capi as ocapi
from tests.gpu_util import assert_fs_equal, assert_stats_equal

pytestmark = pytest.mark.gpu


def test_cfg1_one_million_transactions_reference_settings():
    S, J, n = 200, 2, 1_000_000
    spec = syn.StreamSpec(seed=0xA9B200, n=n, num_services=S, num_jvms=J, span_ms=6 * 3600 * 1000, jitter_ms=40000,
                          zipf_s=1.1, shift_mod=40)
    stream = syn.SynthStream(spec, device=False)
    svc, jvm, ts, el = stream.generate_host()
    lags, thr, infl = (360, 8640), (20.0, 15.0), (0.1, 0.0)
    eng = Engine(EngineConfig(num_services=S, num_jvms=J, lags=lags, thresholds=thr, influences=infl, out_ticks=256,
                              max_tx_per_bucket=1 << 16, max_batch=1 << 17, max_alerts_per_tick=4096))
    orc = ocapi.Pipeline(ocapi.make_config(S, J, 10000, 30, 6, (75.0, 95.0), (0, 1), lags, thr, infl), fast=True)
    # the one-day lag only warms up after 8640 ticks: give 40 keys a resumed history, as a resume file would
    rng = np.random.default_rng(5)
    for key in rng.choice(S * J, 40, replace=False):
        for s in range(3):
            hist = np.round(rng.uniform(50, 400) + rng.normal(0, 3, 8640), 1)
            hist[rng.integers(0, 8640, 20)] = np.nan  # empty windows inside the history
            eng.ring_load(int(key), 1, s, hist)
            orc.ring_load(int(key), 1, s, hist)
    orc.consume(svc, jvm, ts, el)
    st_parts, fs_parts = [], []
    step = 50_000  # ~108 ticks per chunk, below out_ticks
    for a in range(0, n, step):
        eng.process_host(svc[a:a + step], jvm[a:a + step], ts[a:a + step], el[a:a + step])
        st_parts.append(eng.fetch_stats().copy())
        fs_parts.append(eng.fetch_fs().copy())
        eng.release_ticks()
    st, fs = np.concatenate(st_parts), np.concatenate(fs_parts)
    assert orc.ticks > 2000
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())
    assert int(np.abs(fs["signal"]).sum()) > 0
    c = eng.counters()
    assert c.ticks == orc.ticks and c.tx_ingested + c.tx_dropped == n
    eng.close()


def test_cfg3_hundred_million_transactions_sampled_keys_and_invariants():
    import bench
    wl = bench.WORKLOADS["cfg3"]
    K, n = wl["services"], wl["n"]
    span_ms = wl["span_buckets"] * wl["bucket_ms"]
    per_bucket = n // wl["span_buckets"]
    L = wl["lags"][0]
    spec = syn.StreamSpec(seed=wl["seed"], n=n, num_services=K, span_ms=span_ms, shift_mod=97)
    stream = syn.SynthStream(spec, device=True)
    cols = syn.DeviceColumns(n)
    stream.generate_device(cols)
    flags = _capi.F_KEEP_FS | _capi.F_KEEP_STATS | _capi.F_ALERTS | _capi.F_ZSCORE | _capi.F_PERCENTILES
    eng = Engine(EngineConfig(num_services=K, bucket_ms=wl["bucket_ms"], window=1, buffer=0, percentiles=wl["percentiles"],
                              z_series_pct=wl["z_series_pct"], lags=wl["lags"], thresholds=wl["thresholds"],
                              influences=wl["influences"], flags=flags, max_tx_per_bucket=int(per_bucket * 1.5),
                              max_batch=1 << 24, out_ticks=72, max_alerts_per_tick=K))
    buf = syn.DeviceBuffer(L * 3 * K * 8)
    syn.ring_fill_device(buf, L, stream, 77)
    eng.ring_load_all(0, length=L, device_ptr=buf.ptr)
    buf.free()
    res = eng.process_device(cols.columns(), n)
    st, fs = eng.fetch_stats(), eng.fetch_fs()
    svc, jvm, ts, el = cols.to_host()
    cols.free()

    # ---- invariants over ALL keys
    assert res.tx_ingested + res.tx_dropped + res.tx_invalid == n and res.tx_invalid == 0
    labels = ts // wl["bucket_ms"]
    run = np.maximum.accumulate(labels)
    tick_pos = np.flatnonzero(np.concatenate([[True], run[1:] > run[:-1]]))
    assert res.ticks == tick_pos.size and tick_pos.size >= wl["span_buckets"]  # the jitter can open one more bucket
    tick_ts = np.unique(st["tick_ts"])
    assert tick_ts.size == res.ticks - 1  # the first tick fires before any key is seen
    for i in range(1, tick_pos.size):
        closing = labels[tick_pos[i]] - 1  # window = {latest-1, latest}; latest is still empty when the tick fires
        want = int(np.count_nonzero(labels[:tick_pos[i]] == closing))
        rows = st[st["tick_ts"] == (closing * wl["bucket_ms"])]  # edgeTs = (latest - 0 - 1) * bucket_ms
        assert int(rows["count"].sum()) == want, i
        assert int(rows["sum"].sum()) == int(el[:tick_pos[i]][labels[:tick_pos[i]] == closing].sum(dtype=np.int64))

    # ---- a 1 % key sample, record by record, against the oracle
    sample = np.sort(np.random.default_rng(11).choice(K, 100, replace=False))
    sample[:3] = (0, 1, 2)  # the Zipf-hot keys are in
    sample = np.unique(sample)
    keep = np.zeros(K, np.uint8)
    keep[sample] = 1
    oflags = ocapi.F_KEEP_FS | ocapi.F_KEEP_STATS | ocapi.F_ZSCORE | ocapi.F_PERCENTILES
    orc = ocapi.Pipeline(ocapi.make_config(K, 1, wl["bucket_ms"], 1, 0, wl["percentiles"], wl["z_series_pct"], wl["lags"],
                                           wl["thresholds"], wl["influences"], flags=oflags), fast=True)
    orc.set_filter(keep)
    host_stream = syn.SynthStream(spec, device=False)
    for s in range(3):
        vals = syn.ring_fill_host(L, host_stream, 77, s * K + sample)
        for j, key in enumerate(sample):
            orc.ring_load(int(key), 0, s, vals[:, j])
    orc.consume(svc, jvm, ts, el)
    ost, ofs = orc.stats(), orc.fs()
    m_st, m_fs = np.isin(st["key"], sample), np.isin(fs["key"], sample)
    # a sampled key first seen later in the stream has no record before its first transaction on either side
    assert_stats_equal(st[m_st], ost)
    assert_fs_equal(fs[m_fs], ofs)
    assert ofs.size > 50 * sample.size
    eng.close()


def test_cfg5_heavy_skew_top10_services_carry_80_percent():
    """BASELINE configs[4]'s distribution (top 10 services = 80 % of the traffic) on one GPU at a
    reduced length (8 M transactions): the hot keys' 100 k-value runs go through the warp-combined
    path of the accumulator rebuild and the long-run percentile selection; all of them plus a
    sample of cold keys are compared with the oracle, and every tick's totals are checked."""
    K, n, buckets = 10_000, 8_000_000, 12
    spec = syn.StreamSpec(seed=4, n=n, num_services=K, span_ms=buckets * 60_000, hot=(10, 0.8), shift_mod=3)
    stream = syn.SynthStream(spec, device=True)
    cols = syn.DeviceColumns(n)
    stream.generate_device(cols)
    lags, thr, infl = (6,), (1.5,), (0.3,)
    eng = Engine(EngineConfig(num_services=K, bucket_ms=60_000, window=1, buffer=0, percentiles=(50.0, 95.0, 99.0),
                              z_series_pct=(1, 2), lags=lags, thresholds=thr, influences=infl,
                              max_tx_per_bucket=int(n / buckets * 1.5), max_batch=1 << 22, out_ticks=32, max_alerts_per_tick=K))
    res = eng.process_device(cols.columns(), n)
    st, fs = eng.fetch_stats(), eng.fetch_fs()
    svc, jvm, ts, el = cols.to_host()
    cols.free()
    assert res.tx_ingested + res.tx_dropped == n
    hot_share = np.count_nonzero(svc < 10) / n
    assert 0.78 < hot_share < 0.82
    sample = np.unique(np.concatenate([np.arange(10), np.random.default_rng(3).choice(K, 60, replace=False)]))
    keep = np.zeros(K, np.uint8)
    keep[sample] = 1
    orc = ocapi.Pipeline(ocapi.make_config(K, 1, 60_000, 1, 0, (50.0, 95.0, 99.0), (1, 2), lags, thr, infl), fast=True)
    orc.set_filter(keep)
    orc.consume(svc, jvm, ts, el)
    assert_stats_equal(st[np.isin(st["key"], sample)], orc.stats())
    assert_fs_equal(fs[np.isin(fs["key"], sample)], orc.fs())
    assert int(st[st["key"] < 10]["count"].max()) > 40_000
    labels = ts // 60_000
    run = np.maximum.accumulate(labels)
    tick_pos = np.flatnonzero(np.concatenate([[True], run[1:] > run[:-1]]))
    for i in range(1, tick_pos.size):
        closing = labels[tick_pos[i]] - 1
        rows = st[st["tick_ts"] == closing * 60_000]
        assert int(rows["count"].sum()) == int(np.count_nonzero(labels[:tick_pos[i]] == closing))
    eng.close()


@pytest.mark.parametrize("mode", ["exact", "incremental"])
def test_cfg4_shard_size_sampled_keys(mode):
    """One rank's share of BASELINE configs[3] at 8 GPUs: 125 M transactions over 125 k keys, 10 080-slot lists
    (30 GB of rings, pre-filled), p50/p95/p99 + z-score.  A 0.4 % key sample is followed record by record by the
    oracle (ticks stay global: it also sees every tick-triggering transaction); with the exact-order baseline every
    field is bit-identical, with the guarded incremental one signals and values are, means / bounds within 1e-9."""
    import bench
    wl = bench.WORKLOADS["cfg4_shard"]
    K, n = wl["services"], wl["n"]
    span_ms = wl["span_buckets"] * wl["bucket_ms"]
    per_bucket = n // wl["span_buckets"]
    L = wl["lags"][0]
    spec = syn.StreamSpec(seed=wl["seed"], n=n, num_services=K, span_ms=span_ms, shift_mod=97)
    stream = syn.SynthStream(spec, device=True)
    cols = syn.DeviceColumns(n)
    stream.generate_device(cols)
    flags = _capi.F_KEEP_FS | _capi.F_KEEP_STATS | _capi.F_ALERTS | _capi.F_ZSCORE | _capi.F_PERCENTILES
    if mode == "incremental":
        flags |= _capi.F_INCREMENTAL
    eng = Engine(EngineConfig(num_services=K, bucket_ms=wl["bucket_ms"], window=1, buffer=0, percentiles=wl["percentiles"],
                              z_series_pct=wl["z_series_pct"], lags=wl["lags"], thresholds=wl["thresholds"],
                              influences=wl["influences"], flags=flags, max_tx_per_bucket=int(per_bucket * 1.5),
                              max_batch=1 << 25, out_ticks=12, max_alerts_per_tick=K))
    buf = syn.DeviceBuffer(L * 3 * K * 8)
    syn.ring_fill_device(buf, L, stream, 77)
    eng.ring_load_all(0, length=L, device_ptr=buf.ptr)
    buf.free()
    sample = np.unique(np.concatenate([np.arange(4), np.random.default_rng(13).choice(K, 500, replace=False)]))
    st_parts, fs_parts, ticks = [], [], 0
    step = 8 * per_bucket  # out_ticks is small (records of 125 k keys x 60 ticks would be 1.7 GB): fetch every 8 buckets
    ing = 0
    for a in range(0, n, step):
        m = min(step, n - a)
        res = eng.process_device(cols.columns(a), m)
        ticks += res.ticks
        ing += res.tx_ingested + res.tx_dropped + res.tx_invalid
        st, fs = eng.fetch_stats(), eng.fetch_fs()
        st_parts.append(st[np.isin(st["key"], sample)].copy())
        fs_parts.append(fs[np.isin(fs["key"], sample)].copy())
        eng.release_ticks()
    assert ing == n and ticks >= wl["span_buckets"]
    st, fs = np.concatenate(st_parts), np.concatenate(fs_parts)
    svc, jvm, ts, el = cols.to_host()
    cols.free()
    lab = ts // wl["bucket_ms"]
    run = np.maximum.accumulate(np.r_[0, lab])
    keep_rows = np.isin(svc, sample) | (run[1:] > run[:-1])
    keep = np.zeros(K, np.uint8)
    keep[sample] = 1
    oflags = ocapi.F_KEEP_FS | ocapi.F_KEEP_STATS | ocapi.F_ZSCORE | ocapi.F_PERCENTILES
    orc = ocapi.Pipeline(ocapi.make_config(K, 1, wl["bucket_ms"], 1, 0, wl["percentiles"], wl["z_series_pct"], wl["lags"],
                                           wl["thresholds"], wl["influences"], flags=oflags), fast=True)
    orc.set_filter(keep)
    host_stream = syn.SynthStream(spec, device=False)
    for s in range(3):
        vals = syn.ring_fill_host(L, host_stream, 77, s * K + sample)
        for j, key in enumerate(sample):
            orc.ring_load(int(key), 0, s, vals[:, j])
    orc.consume(svc[keep_rows], jvm[keep_rows], ts[keep_rows], el[keep_rows])
    assert orc.ticks == ticks
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs(), exact=(mode == "exact"))
    assert fs.size > 40 * sample.size
    eng.close()
