"""The stats-only path (BASELINE config 2: count / mean / stddev / min / max, no percentiles, no z-score) handles a
whole chunk in two launches - one keyed reduce into per-segment planes, one kernel that walks the chunk's ticks per
key (k1_chunk.cu).  Its records must equal the oracle's and those of the per-tick path (APM_STATS_CHUNK=0) bit for
bit: same ticks, same windows, same first-seen behaviour, also when ticks fall inside a CTA's stretch, when buckets
are skipped, when keys appear in the middle of a chunk and when durations leave the shared-memory table's range -
with the direct-mapped table (every key fits shared memory) and with the hashed one."""
import os

import numpy as np
import pytest

from apmbackend_b200 import Engine, EngineConfig, _capi
from oracle import capi as ocapi
from tests.gpu_util import assert_stats_equal
from tests.streams import make_stream

pytestmark = pytest.mark.gpu


def _engine(S, J, bucket_ms, window, buffer, env, **kw):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)  # the knobs are read when the context is created
    try:
        return Engine(EngineConfig(num_services=S, num_jvms=J, bucket_ms=bucket_ms, window=window, buffer=buffer, percentiles=(), lags=(),
                                   thresholds=(), influences=(), flags=_capi.F_KEEP_STATS, **kw))
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


CHUNKED = {"APM_STATS_CHUNK": "1", "APM_STATS_CHUNK_MIN_N": "1", "APM_STATS_CHUNK_MIN_SEG": "1"}
PER_TICK = {"APM_STATS_CHUNK": "0"}


def _run(eng, cols, calls):
    n = cols[0].size
    out = []
    bounds = np.linspace(0, n, calls + 1).astype(int)
    for a, b in zip(bounds[:-1], bounds[1:]):
        eng.process_host(*(x[a:b] for x in cols))
        out.append(eng.fetch_stats().copy())
        eng.release_ticks()
    c = eng.counters()
    eng.close()
    return np.concatenate(out), (c.tx_ingested, c.tx_dropped, c.tx_invalid, c.ticks, c.keys_seen)


@pytest.mark.parametrize("S,J,n,window,buffer,span,hi,late,gap,max_batch,calls", [
    (50, 2, 60000, 1, 0, 40, 60000, 0.02, None, 1 << 14, 1),          # tumbling window, several chunks per call
    (300, 1, 90000, 3, 1, 60, 60000, 0.05, (20, 26), 1 << 15, 3),     # sliding window, a gap (several slots evicted at once)
    (20000, 1, 400000, 2, 1, 12, 60000, 0.02, None, 1 << 17, 1),      # more keys than a CTA's table holds
    (40, 1, 50000, 30, 6, 300, 5_000_000, 0.1, None, 1 << 13, 2),     # the reference's window; durations above 2^16 ms
], ids=["tumbling", "sliding-gap", "many-keys", "reference-window-wide-values"])
def test_whole_chunk_path_equals_the_oracle_and_the_per_tick_path(S, J, n, window, buffer, span, hi, late, gap, max_batch, calls):
    bucket_ms = 10000
    svc, jvm, ts, el = make_stream(97, n, S, J, bucket_ms, span, elapsed_hi=hi, late_frac=late, late_ms=45000, gap_at=gap)
    if hi > 65536:
        el = el.copy()
        el[::5] = np.minimum(el[::5].astype(np.int64) * 700, hi).astype(np.int32)
    svc, el = svc.copy(), el.copy()
    svc[::1013] = S + 1      # invalid rows
    el[7::2027] = -3
    # one key is named for the first time in the middle of the stream: no record before that, one ever after
    late_key = S - 1
    first_at = n // 2
    svc[:first_at][svc[:first_at] == late_key] = 0
    svc[first_at] = late_key
    jvm = jvm.copy(); jvm[first_at] = 0
    cols = (svc, jvm, ts, el)
    kw = dict(max_batch=max_batch, out_ticks=span + 8)
    got, cnt = _run(_engine(S, J, bucket_ms, window, buffer, CHUNKED, **kw), cols, calls)
    want, wcnt = _run(_engine(S, J, bucket_ms, window, buffer, PER_TICK, **kw), cols, calls)
    assert got.size == want.size and got.tobytes() == want.tobytes()
    assert cnt == wcnt
    # up to 11 584 keys every key has its own shared-memory entry (no tags, no probing); the hashed table gives the same
    hashed, hcnt = _run(_engine(S, J, bucket_ms, window, buffer, dict(CHUNKED, APM_STATS_CHUNK_DIRECT="0"), **kw), cols, calls)
    assert hashed.tobytes() == want.tobytes() and hcnt == wcnt
    orc = ocapi.Pipeline(ocapi.make_config(S, J, bucket_ms, window, buffer, (), (0, 1), (), (), (), flags=ocapi.F_KEEP_STATS), fast=True)
    orc.consume(*cols)
    assert_stats_equal(got, orc.stats())
    assert cnt[3] == orc.ticks
    rows = got[got["key"] == late_key]
    assert rows.size > 0 and rows["tick_ts"].min() > ts[first_at] - (buffer + 2) * bucket_ms


def test_whole_chunk_path_is_the_default_for_large_stats_only_chunks():
    """The default thresholds: a chunk of a few hundred thousand transactions with long segments goes the whole-chunk
    way (two launches per chunk instead of three per tick), a small one keeps the per-tick kernels."""
    S, bucket_ms = 2000, 60000
    svc, jvm, ts, el = make_stream(98, 600000, S, 1, bucket_ms, 20, late_frac=0.01)
    big = Engine(EngineConfig(num_services=S, bucket_ms=bucket_ms, window=1, buffer=0, percentiles=(), lags=(), thresholds=(), influences=(),
                              flags=_capi.F_KEEP_STATS, max_batch=1 << 20, out_ticks=32))
    big.process_host(svc, jvm, ts, el)
    launches_big = big.counters().kernel_launches
    st = big.fetch_stats().copy()
    big.close()
    orc = ocapi.Pipeline(ocapi.make_config(S, 1, bucket_ms, 1, 0, (), (0, 1), (), (), (), flags=ocapi.F_KEEP_STATS), fast=True)
    orc.consume(svc, jvm, ts, el)
    assert_stats_equal(st, orc.stats())
    assert launches_big <= 8, launches_big  # tick finder (3) + reduce + ticks
    small = _engine(S, 1, bucket_ms, 1, 0, PER_TICK, max_batch=1 << 20, out_ticks=32)
    small.process_host(svc, jvm, ts, el)
    assert small.counters().kernel_launches > 3 * orc.ticks
    small.close()


def test_cfg2_size_100m_transactions_sampled_keys():
    """BASELINE config 2 at size: 100 M transactions, 10 k keys, 60 one-minute buckets, stats only.  Whole-stream
    invariants plus the records of a 1 % key sample against the oracle (which is fed that sample's transactions and
    every tick-triggering one)."""
    from apmbackend_b200 import synthetic as syn
    n, K, span_buckets, bucket_ms = 100_000_000, 10_000, 60, 60_000
    spec = syn.StreamSpec(seed=1, n=n, num_services=K, num_jvms=1, span_ms=span_buckets * bucket_ms)
    stream = syn.SynthStream(spec, device=True)
    dc = syn.DeviceColumns(n)
    stream.generate_device(dc)
    eng = Engine(EngineConfig(num_services=K, bucket_ms=bucket_ms, window=1, buffer=0, percentiles=(), lags=(), thresholds=(), influences=(),
                              flags=_capi.F_KEEP_STATS, max_batch=1 << 25, out_ticks=span_buckets + 4))
    eng.process_device(dc.columns(), n)
    st = eng.fetch_stats().copy()
    cnt = eng.counters()
    eng.close()
    svc, jvm, ts, el = dc.to_host()
    dc.free(); stream.free()
    assert cnt.tx_ingested + cnt.tx_dropped == n and cnt.ticks >= span_buckets
    # a tumbling window counts a transaction at most once (late arrivals and the stream's last bucket: never)
    assert 0.95 * n < int(st["count"].sum()) <= int(cnt.tx_ingested)
    keys = np.arange(37, K, 100)
    lab = ts // bucket_ms
    run = np.maximum.accumulate(np.r_[0, lab])
    trig = run[1:] > run[:-1]
    keep = np.isin(svc, keys) | trig
    orc = ocapi.Pipeline(ocapi.make_config(K, 1, bucket_ms, 1, 0, (), (0, 1), (), (), (), flags=ocapi.F_KEEP_STATS), fast=True)
    filt = np.zeros(K, np.uint8)
    filt[keys] = 1
    orc.set_filter(filt)
    orc.consume(svc[keep], jvm[keep], ts[keep], el[keep])
    want = orc.stats()
    want = want[np.isin(want["key"], keys)]
    got = st[np.isin(st["key"], keys)]
    assert got.size == want.size > 0
    assert_stats_equal(got, want)
