"""Key-sharded runs on the GPU: the global stream is binned by owner on the host (apm_partition_host), every shard
runs in its own context with the GLOBAL ticks handed over as marks (apm_set_tick_marks), and the union of the
shards' records must equal, bit for bit, what one unsharded context - and the oracle - produce over the whole
stream.  No statistic has a cross-key term (stream_calc_stats.js:159-202, stream_calc_z_score.js:197-308), but a
tick is triggered by ANY transaction (:348), so a shard must tick while it is silent and must not tick late when
its own first transaction of a bucket arrives after another shard's.  All contexts share device 0 here; the NCCL
gather over several GPUs is covered by test_gpu_gather_nccl.py and by bench.py's sharded_parity leg."""
import numpy as np
import pytest

from apmbackend_b200 import Engine, _capi, sharding
from apmbackend_b200._capi import ApmError
from oracle import capi as ocapi
from tests.gpu_util import assert_fs_equal, assert_stats_equal, engine_config, oracle_config
from tests.streams import make_stream

pytestmark = pytest.mark.gpu

BASE = dict(S=40, J=2, bucket_ms=10000, window=3, buffer=1, lags=(6, 15), thr=(1.5, 2.5), infl=(0.3, 0.0),
            pcts=(75.0, 95.0))


def _local_ids(own, K, nranks):
    local = np.zeros(K, np.uint32)
    counts = []
    for r in range(nranks):
        idx = np.flatnonzero(own == r)
        local[idx] = np.arange(idx.size, dtype=np.uint32)
        counts.append(idx.size)
    return local, counts


def _run_sharded(c, cols, nranks, owner_table=None, calls=1, via="host", max_batch=None, dense_ids=True, flags=None):
    """Returns (stat records, fs records, alert records) of all shards with keys mapped back to global ids."""
    K = c["S"] * c["J"]
    own = sharding.owner_of_keys(np.arange(K), nranks) if owner_table is None else np.asarray(owner_table)
    local, counts = _local_ids(own, K, nranks)
    glob = [np.flatnonzero(own == r) for r in range(nranks)]  # local id -> global key
    engines = []
    for r in range(nranks):
        cc = dict(c)
        if dense_ids:
            cc["S"], cc["J"] = max(counts[r], 1), 1
        if max_batch:
            cc["max_batch"] = max_batch
        kw = {} if flags is None else {"flags": flags}
        engines.append(Engine(engine_config(cc, **kw)))
    n = cols[0].size
    bounds = np.linspace(0, n, calls + 1).astype(int)
    latest = 0
    st_all, fs_all, al_all = [], [], []
    for a, b in zip(bounds[:-1], bounds[1:]):
        parts, latest_next = sharding.partition_native(*(x[a:b] for x in cols), c["bucket_ms"], c["S"], c["J"], nranks,
                                                       latest=latest, owner_table=owner_table,
                                                       local_id=local if dense_ids else None)
        for r, eng in enumerate(engines):
            (svc, jvm, ts, el), marks = parts[r]
            eng.set_tick_marks(marks)
            if via == "host":
                eng.process_host(svc, jvm, ts, el)
            elif via == "packed":
                base = int(ts.min()) if ts.size else 0
                eng.process_host_packed((jvm << 16 | svc).astype(np.uint32), (ts - base).astype(np.uint32), el.astype(np.uint32), base)
            else:
                from apmbackend_b200 import synthetic as syn
                dc = syn.DeviceColumns(max(svc.size, 1))
                for buf, arr in zip(dc.bufs, (svc, jvm, ts, el)):
                    if arr.size:
                        buf.upload(arr)
                eng.process_device(dc.columns(), svc.size)
                eng.sync()
                dc.free()
            assert eng.latest == latest_next  # every shard follows the global latestBucket
            st, fs, al = eng.fetch_stats().copy(), eng.fetch_fs().copy(), eng.fetch_alerts().copy()
            if dense_ids:
                for rec in (st, fs, al):
                    rec["key"] = glob[r][rec["key"]]
            st_all.append(st); fs_all.append(fs); al_all.append(al)
            eng.release_ticks()
        latest = latest_next
    ticks = [e.counters().ticks for e in engines]
    for e in engines:
        e.close()
    return np.concatenate(st_all), np.concatenate(fs_all), np.concatenate(al_all), ticks


def _unsharded(c, cols):
    eng = Engine(engine_config(c))
    eng.process_host(*cols)
    out = eng.fetch_stats().copy(), eng.fetch_fs().copy(), eng.fetch_alerts().copy(), eng.counters().ticks
    eng.close()
    return out


def _sorted_bytes(a):
    return a[np.lexsort((a["key"], a["lag"], a["tick_ts"]))].tobytes()


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_sharded_union_equals_the_unsharded_run_and_the_oracle(nranks):
    c = dict(BASE)
    # few transactions per bucket and a gap: most shards are silent in many buckets, all are silent in some
    cols = make_stream(61, 6000, c["S"], c["J"], c["bucket_ms"], 150, late_frac=0.08, gap_at=(60, 70), shift_services=(0, 1, 2, 3, 5))
    st, fs, al, ticks = _run_sharded(c, cols, nranks)
    ust, ufs, ual, uticks = _unsharded(c, cols)
    assert all(t == uticks for t in ticks)
    assert_stats_equal(st, ust)
    assert_fs_equal(fs, ufs)
    assert ual.size > 0 and _sorted_bytes(al) == _sorted_bytes(ual)
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    orc.consume(*cols)
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())


def test_without_global_ticks_a_shard_would_differ():
    """The property the marks exist for: left to find ticks in its own sub-stream, a shard ticks late (its first
    transaction of a bucket comes after the global trigger) or not at all (silent buckets) - the records differ."""
    c = dict(BASE)
    cols = make_stream(61, 6000, c["S"], c["J"], c["bucket_ms"], 150, late_frac=0.08, gap_at=(60, 70))
    parts, _ = sharding.partition_native(*cols, c["bucket_ms"], c["S"], c["J"], 8)
    ust, _, _, uticks = _unsharded(c, cols)
    eng = Engine(engine_config(c))
    eng.process_host(*parts[3][0])  # no marks: the device finds this shard's own ticks
    st = eng.fetch_stats().copy()
    assert eng.counters().ticks < uticks or st.tobytes() != ust[np.isin(ust["key"], st["key"])].tobytes()
    eng.close()


@pytest.mark.parametrize("via", ["host", "packed", "device"])
def test_marks_work_through_every_process_entry_point_and_across_chunks(via):
    c = dict(BASE, S=30, J=1)
    cols = make_stream(62, 9000, c["S"], c["J"], c["bucket_ms"], 90, late_frac=0.05, shift_services=(0, 2))
    # max_batch far below a call: marks land on chunk boundaries, inside chunks and behind the last transaction
    st, fs, al, _ = _run_sharded(c, cols, 3, calls=4, via=via, max_batch=256)
    ust, ufs, ual, _ = _unsharded(c, cols)
    assert_stats_equal(st, ust)
    assert_fs_equal(fs, ufs)
    assert _sorted_bytes(al) == _sorted_bytes(ual)


def test_owner_table_placement_and_global_ids():
    c = dict(BASE)
    cols = make_stream(63, 5000, c["S"], c["J"], c["bucket_ms"], 100, shift_services=(0, 1))
    K = c["S"] * c["J"]
    key = cols[1].astype(np.int64) * c["S"] + cols[0]
    table = sharding.balanced_owner_table(np.bincount(key, minlength=K), 4, key_cost=20)
    ust, ufs, _, _ = _unsharded(c, cols)
    for dense in (True, False):  # dense per-owner ids, or every context sized for the global key space
        st, fs, _, _ = _run_sharded(c, cols, 4, owner_table=table, dense_ids=dense)
        assert_stats_equal(st, ust)
        assert_fs_equal(fs, ufs)


def test_a_shard_without_any_transaction_in_a_call_still_ticks():
    c = dict(BASE, S=8, J=1)
    cols = make_stream(64, 800, c["S"], c["J"], c["bucket_ms"], 80)
    K = c["S"]
    table = np.zeros(K, np.uint32)
    table[K - 1] = 1  # rank 1 owns one rare key: whole calls go by without a transaction for it
    st, fs, _, ticks = _run_sharded(c, cols, 2, owner_table=table, calls=16)
    ust, ufs, _, uticks = _unsharded(c, cols)
    assert ticks == [uticks, uticks]
    assert_stats_equal(st, ust)
    assert_fs_equal(fs, ufs)


def test_bad_marks_are_refused():
    c = dict(BASE, S=4, J=1)
    svc, jvm, ts, el = make_stream(65, 200, 4, 1, c["bucket_ms"], 10)
    lab0 = int(ts[0] // c["bucket_ms"])
    eng = Engine(engine_config(c))
    with pytest.raises(ApmError):
        eng.set_tick_marks([(5, lab0 + 1), (3, lab0 + 2)])      # positions go back
    with pytest.raises(ApmError):
        eng.set_tick_marks([(0, lab0 + 2), (3, lab0 + 2)])      # labels must grow
    eng.set_tick_marks([(0, lab0), (svc.size + 1, lab0 + 1)])   # a mark beyond the call's transactions
    with pytest.raises(ApmError) as ei:
        eng.process_host(svc, jvm, ts, el)
    assert ei.value.status == _capi.E_INVALID
    # the marks were consumed: the next call finds its ticks on the device again
    eng2 = Engine(engine_config(c))
    eng2.process_host(svc, jvm, ts, el)
    assert eng2.counters().ticks > 0
    eng.close(); eng2.close()
