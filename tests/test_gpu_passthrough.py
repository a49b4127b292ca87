"""the time-ordered transaction passthrough (stream_calc_stats.js:133, 136-155 over
binary_heap.js) on the device.  Per tick the device must release exactly the transactions the
reference's heap pops (equal multiset) in non-decreasing endTs order."""
import numpy as np
import pytest

from apmbackend_b200 import Engine, EngineConfig, _capi
from apmbackend_b200.binary_heap import BinaryHeap
from tests.streams import make_stream

pytestmark = pytest.mark.gpu


def _reference_heap(ts, bucket_ms, buffer):
    """consumeMsg's heap traffic: pop everything <= edgeTs at a tick, then push the transaction."""
    heap = BinaryHeap(lambda e: e[0])
    latest, per_tick = 0, []
    for i, t in enumerate(ts.tolist()):
        label = t // bucket_ms
        if label > latest:
            latest = label
            edge = (latest - buffer - 1) * bucket_ms
            per_tick.append(heap.popAllLessOrEqualToScore(edge))
        heap.push((t, i))
    return per_tick


@pytest.mark.parametrize("late_frac,chunk", [(0.02, 700), (0.3, 5000)])
def test_passthrough_equals_the_reference_heap(late_frac, chunk):
    S, J, bucket_ms, buffer = 6, 2, 10000, 6
    svc, jvm, ts, el = make_stream(71, 6000, S, J, bucket_ms, 120, late_frac=late_frac, late_ms=150000)
    flags = _capi.F_KEEP_STATS | _capi.F_PASSTHROUGH
    eng = Engine(EngineConfig(num_services=S, num_jvms=J, percentiles=(), lags=(), thresholds=(), influences=(), flags=flags,
                              max_tx_per_bucket=1 << 14, max_batch=1 << 14, out_ticks=256))
    want = _reference_heap(ts, bucket_ms, buffer)
    got_ts, got_seq, got_counts = [], [], []
    for a in range(0, ts.size, chunk):
        eng.process_host(svc[a:a + chunk], jvm[a:a + chunk], ts[a:a + chunk], el[a:a + chunk])
        t, s, c = eng.fetch_passthrough()
        got_ts.append(t.copy()); got_seq.append(s.copy()); got_counts.append(c.copy())
        eng.release_ticks()
    got_ts, got_seq, got_counts = np.concatenate(got_ts), np.concatenate(got_seq), np.concatenate(got_counts)
    assert got_counts.size == len(want)
    assert [int(c) for c in got_counts] == [len(w) for w in want]
    off = 0
    for w in want:
        n = len(w)
        seg_ts, seg_seq = got_ts[off:off + n], got_seq[off:off + n]
        assert np.all(np.diff(seg_ts) >= 0)                                   # non-decreasing endTs
        assert sorted(seg_seq.tolist()) == sorted(i for _, i in w)            # equal multiset
        assert seg_ts.tolist() == [t for t, _ in w]                           # the heap pops in score order, too
        assert np.array_equal(ts[seg_seq], seg_ts)                            # seq really names that transaction
        off += n
    assert off == got_ts.size and off > 4000
    eng.close()


def test_passthrough_with_the_commit_tick_protocol_and_invalid_rows():
    S, bucket_ms, buffer = 4, 10000, 6
    svc, jvm, ts, el = make_stream(72, 1500, S, 1, bucket_ms, 60)
    svc = svc.copy(); svc[100] = 99  # invalid row: takes a sequence number, never enters the buffer
    flags = _capi.F_KEEP_STATS | _capi.F_PASSTHROUGH
    eng = Engine(EngineConfig(num_services=S, percentiles=(), lags=(), thresholds=(), influences=(), flags=flags,
                              max_tx_per_bucket=1 << 12, max_batch=1 << 12, out_ticks=128))
    labels = ts // bucket_ms
    latest, start = 0, 0
    for i in range(ts.size):
        if labels[i] > latest:
            eng.add_batch(svc[start:i], jvm[start:i], ts[start:i], el[start:i])
            start, latest = i, int(labels[i])
            eng.tick(latest)
    t, s, c = eng.fetch_passthrough()
    assert 100 not in s.tolist() and np.array_equal(ts[s], t)
    edge = (latest - buffer - 1) * bucket_ms
    due = np.flatnonzero((ts[:start] <= edge) & (np.arange(start) != 100))
    assert sorted(s.tolist()) == due.tolist()
    eng.close()


def test_passthrough_survives_save_and_load(tmp_path):
    """The reference writes `minHeap` into the stats resume file and re-pushes it on resume
    (stream_calc_stats.js:40-52, 54-61): a context restored from a state image must release the transactions that
    were still waiting, under their old sequence numbers, exactly as the context that kept running does."""
    S, J, bucket_ms = 6, 2, 10000
    svc, jvm, ts, el = make_stream(73, 5000, S, J, bucket_ms, 100, late_frac=0.1, late_ms=120000)
    flags = _capi.F_KEEP_STATS | _capi.F_PASSTHROUGH
    mk = lambda: Engine(EngineConfig(num_services=S, num_jvms=J, percentiles=(), lags=(), thresholds=(), influences=(),  # noqa: E731
                                     flags=flags, max_tx_per_bucket=1 << 14, max_batch=1 << 14, out_ticks=256))
    half = 2600
    a = mk()
    a.process_host(svc[:half], jvm[:half], ts[:half], el[:half])
    first = a.fetch_passthrough()
    a.release_ticks()
    path = str(tmp_path / "pt.state")
    a.state_save(path)
    b = mk()
    b.state_load(path)
    outs = []
    for eng in (a, b):
        eng.process_host(svc[half:], jvm[half:], ts[half:], el[half:])
        t, s, c = eng.fetch_passthrough()
        outs.append((t.copy(), s.copy(), c.copy(), eng.fetch_stats().copy()))
        eng.close()
    for x, y in zip(outs[0], outs[1]):
        assert x.tobytes() == y.tobytes()
    # and the two halves together are what one uninterrupted run releases
    want = _reference_heap(ts, bucket_ms, 6)
    got = np.concatenate([first[1], outs[1][1]])
    assert sorted(got.tolist()) == sorted(i for w in want for _, i in w)
    assert (outs[1][1] < half).any(), "some transactions of the first half must have been waiting across the checkpoint"


def test_set_latest_gives_the_passthrough_its_time_base():
    """After apm_set_latest (resume from the reference's JSON file) data may arrive before the first tick."""
    S, bucket_ms = 4, 10000
    svc, jvm, ts, el = make_stream(74, 400, S, 1, bucket_ms, 12)
    flags = _capi.F_KEEP_STATS | _capi.F_PASSTHROUGH
    eng = Engine(EngineConfig(num_services=S, percentiles=(), lags=(), thresholds=(), influences=(), flags=flags,
                              max_tx_per_bucket=1 << 12, max_batch=1 << 12, out_ticks=64))
    eng.set_latest(int(ts[0] // bucket_ms))
    eng.process_host(svc, jvm, ts, el)  # used to fail: "no tick has set the time base yet"
    t, s, c = eng.fetch_passthrough()
    assert t.size > 0 and np.all(np.diff(t[:c[0]]) >= 0)
    eng.close()
