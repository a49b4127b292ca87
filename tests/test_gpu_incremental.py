"""APM_F_INCREMENTAL: the guarded incremental baseline.  Signals, quantised values and the lag
lists themselves must stay bit-identical to the reference (oracle); the rolling mean and the
bounds may differ by rounding, within 1e-9 relative (BASELINE.json's tolerance for fp64 means)."""
import numpy as np
import pytest

from apmbackend_b200 import Engine, EngineConfig, _capi
from apmbackend_b200 import synthetic as syn
from oracle import capi as ocapi
from tests.gpu_util import assert_fs_equal, assert_stats_equal, engine_config, oracle_config, run_both
from tests.streams import make_stream
from tests.test_oracle_twins import CASES

pytestmark = pytest.mark.gpu
FLAGS = (_capi.F_PERCENTILES | _capi.F_ZSCORE | _capi.F_KEEP_STATS | _capi.F_KEEP_FS | _capi.F_ALERTS | _capi.F_INCREMENTAL)


@pytest.mark.parametrize("c", CASES, ids=lambda c: f"seed{c['seed']}")
def test_incremental_matches_oracle(c):
    cols = make_stream(c["seed"], c["n"], c["S"], c["J"], c["bucket_ms"], c["span"], gap_at=c["gap"], shift_services=(0, 2))
    eng, orc, st, fs = run_both(c, cols, flags=FLAGS)
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs(), exact=False)          # signals / values exact, mean and bounds to 1e-9
    assert int(np.abs(fs["signal"]).sum()) > 0
    for key in range(min(4, c["S"])):                   # the lists themselves are bit-identical
        for li in range(len(c["lags"])):
            for s in range(3):
                assert eng.ring_read(key, li, s).tobytes() == orc.ring_read(key, li, s).tobytes()
    eng.close()


def test_guard_band_falls_back_to_the_exact_sum():
    """Values sitting exactly on the strict `>` boundary: list of 100.0 (sqrt-mean 10, threshold 20)
    and new values 300.0 / 300.1 / -100.0 / -100.1 as in KAT-3 - decided inside the guard band."""
    c = dict(S=2, J=1, bucket_ms=10000, window=1, buffer=0, pcts=(75.0, 95.0), lags=(360,), thr=(20.0,), infl=(0.1,))
    eng = Engine(engine_config(c, flags=FLAGS))
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    for s in range(3):
        eng.ring_load(0, 0, s, np.full(360, 100.0))
        orc.ring_load(0, 0, s, np.full(360, 100.0))
    t0 = 1578409200000
    # one transaction per bucket for key 0 with durations on / around the boundary (avg == p75 == p95 == the duration)
    durs = [300, 301, 100, 300, 299, 300, 301, 300]
    ts = np.array([t0 + i * 10000 for i in range(len(durs) + 2)], np.int64)
    el = np.array(durs + [100, 100], np.int32)
    z = np.zeros(ts.size, np.uint32)
    orc.consume(z, z, ts, el)
    eng.process_host(z, z, ts, el)
    fs = eng.fetch_fs()
    assert_fs_equal(fs, orc.fs(), exact=False)
    assert eng.counters().inc_fallbacks > 0            # the boundary cases went through the exact path
    assert eng.ring_read(0, 0, 0).tobytes() == orc.ring_read(0, 0, 0).tobytes()
    eng.close()


def test_incremental_cfg3_shape_with_loaded_history():
    """10 M transactions in cfg3's shape with a pre-filled 10 080-slot history (bulk load -> the
    running statistics are rebuilt on the device), a key sample against the oracle."""
    K, n, buckets, L = 10_000, 10_000_000, 12, 10080
    spec = syn.StreamSpec(seed=2, n=n, num_services=K, span_ms=buckets * 60_000, shift_mod=97)
    stream = syn.SynthStream(spec, device=True)
    cols = syn.DeviceColumns(n)
    stream.generate_device(cols)
    kw = dict(num_services=K, bucket_ms=60_000, window=1, buffer=0, percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2),
              lags=(L,), thresholds=(15.0,), influences=(0.0,), max_tx_per_bucket=int(n / buckets * 1.5), max_batch=1 << 22,
              out_ticks=32, max_alerts_per_tick=K)
    eng = Engine(EngineConfig(flags=FLAGS, **kw))
    buf = syn.DeviceBuffer(L * 3 * K * 8)
    syn.ring_fill_device(buf, L, stream, 77)
    eng.ring_load_all(0, length=L, device_ptr=buf.ptr)
    buf.free()
    eng.process_device(cols.columns(), n)
    fs = eng.fetch_fs()
    svc, jvm, ts, el = cols.to_host()
    cols.free()
    sample = np.unique(np.concatenate([np.arange(5), np.random.default_rng(9).choice(K, 60, replace=False)]))
    keep = np.zeros(K, np.uint8)
    keep[sample] = 1
    orc = ocapi.Pipeline(ocapi.make_config(K, 1, 60_000, 1, 0, (50.0, 95.0, 99.0), (1, 2), (L,), (15.0,), (0.0,)), fast=True)
    orc.set_filter(keep)
    host_stream = syn.SynthStream(spec, device=False)
    for s in range(3):
        vals = syn.ring_fill_host(L, host_stream, 77, s * K + sample)
        for j, key in enumerate(sample):
            orc.ring_load(int(key), 0, s, vals[:, j])
    orc.consume(svc, jvm, ts, el)
    assert_fs_equal(fs[np.isin(fs["key"], sample)], orc.fs(), exact=False)
    for key in sample[:6]:
        for s in range(3):
            assert eng.ring_read(int(key), 0, s).tobytes() == orc.ring_read(int(key), 0, s).tobytes()
    eng.close()
