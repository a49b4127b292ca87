"""Helpers shared by the GPU parity tests: run the CUDA engine and the CPU oracle on the
same columns and compare their binary records."""
import numpy as np

from apmbackend_b200 import Engine, EngineConfig
from apmbackend_b200 import _capi
from oracle import capi as ocapi


def engine_config(c, **kw):
    return EngineConfig(num_services=c["S"], num_jvms=c["J"], bucket_ms=c["bucket_ms"], window=c["window"],
                        buffer=c["buffer"], percentiles=c["pcts"], z_series_pct=c.get("zpct", (0, 1)),
                        lags=c["lags"], thresholds=c["thr"], influences=c["infl"],
                        max_tx_per_bucket=c.get("log_cap", 1 << 16), max_batch=c.get("max_batch", 1 << 16),
                        out_ticks=c.get("out_ticks", 1024), max_alerts_per_tick=c.get("max_alerts", 4096), **kw)


def oracle_config(c):
    return ocapi.make_config(c["S"], c["J"], c["bucket_ms"], c["window"], c["buffer"], c["pcts"],
                             c.get("zpct", (0, 1)), c["lags"], c["thr"], c["infl"])


def sort_records(a, fs=False):
    """Oracle emits in first-seen order, the engine key-ascending: compare on a common order."""
    if fs:
        # within a tick the engine is lag-major then key; the oracle key-major then lag
        order = np.lexsort((a["lag"], a["key"], a["tick_ts"]))
    else:
        order = np.lexsort((a["key"], a["tick_ts"]))
    return a[order]


def assert_stats_equal(got, want):
    got, want = sort_records(got), sort_records(want)
    assert got.size == want.size, (got.size, want.size)
    for f in ("tick_ts", "key", "flags", "count", "sum", "min", "max"):
        assert np.array_equal(got[f], want[f]), f
    for f in ("tpm", "avg", "pct"):  # single correctly rounded operations on exact integers: bit-exact
        assert np.array_equal(got[f].view(np.uint64), want[f].view(np.uint64)), f
    # extension stddev: same operation sequence on both sides; tolerance 1e-9 relative as the spec allows
    a, b = got["stddev"], want["stddev"]
    assert np.array_equal(np.isnan(a), np.isnan(b))
    m = ~np.isnan(a)
    assert np.all(np.abs(a[m] - b[m]) <= 1e-9 * np.maximum(np.abs(b[m]), 1e-300))


def assert_fs_equal(got, want, exact=True):
    got, want = sort_records(got, True), sort_records(want, True)
    assert got.size == want.size, (got.size, want.size)
    for f in ("tick_ts", "key", "lag", "signal", "flags"):
        assert np.array_equal(got[f], want[f]), f
    for f in ("tpm", "val"):  # quantised inter-stage values: bit-exact
        assert np.array_equal(got[f].view(np.uint64), want[f].view(np.uint64)), f
    for f in ("avg", "lb", "ub"):
        a, b = got[f], want[f]
        assert np.array_equal(np.isnan(a), np.isnan(b)), f
        if exact:  # exact-order summation: identical bits
            assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), f
        else:  # 1e-9 relative to the size of the rolling mean (a bound close to zero is compared on that scale, too)
            m = ~np.isnan(a)
            scale = np.maximum(np.abs(want["avg"]), np.abs(b))
            assert np.all(np.abs(a[m] - b[m]) <= 1e-9 * scale[m]), f


def run_both(c, cols, chunks=None, via="host", **engine_kw):
    """Feeds the same columns to the engine (optionally cut in chunks) and the oracle."""
    svc, jvm, ts, el = cols
    eng = Engine(engine_config(c, **engine_kw))
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    orc.consume(svc, jvm, ts, el)
    n = svc.size
    bounds = [0, n] if not chunks else sorted(set([0, n] + [min(n, x) for x in chunks]))
    st_parts, fs_parts = [], []
    for a, b in zip(bounds[:-1], bounds[1:]):
        eng.process_host(svc[a:b], jvm[a:b], ts[a:b], el[a:b])
        st_parts.append(eng.fetch_stats().copy())
        if c["lags"]:
            fs_parts.append(eng.fetch_fs().copy())
        eng.release_ticks()
    st = np.concatenate(st_parts) if st_parts else np.zeros(0, _capi.STAT_DTYPE)
    fs = np.concatenate(fs_parts) if fs_parts else np.zeros(0, _capi.FS_DTYPE)
    return eng, orc, st, fs
