"""Multi-rank host logic on CPU (gloo, world_size 2): key-hash partition, global ticks and the
alert gather.  The per-rank arithmetic is stood in for by the oracle here - the GPU ranks are
exercised by bench.py --gpus N on the box; what is checked is that sharded results, gathered,
equal the unsharded run."""
import os
import socket

import numpy as np
import pytest

from apmbackend_b200 import _capi, sharding
from oracle import capi as ocapi
from tests.streams import make_stream

CFG = dict(S=12, J=2, bucket_ms=10000, window=3, buffer=1, lags=(6, 15), thr=(1.5, 2.5), infl=(0.3, 0.0),
           pcts=(75.0, 95.0))


def _pipeline():
    return ocapi.Pipeline(ocapi.make_config(CFG["S"], CFG["J"], CFG["bucket_ms"], CFG["window"], CFG["buffer"],
                                            CFG["pcts"], (0, 1), CFG["lags"], CFG["thr"], CFG["infl"]), fast=True)


def _stream():
    return make_stream(31, 6000, CFG["S"], CFG["J"], CFG["bucket_ms"], 120, shift_services=(0, 1, 2, 3))


def test_owner_function_matches_the_c_abi():
    lib = _capi.lib()
    keys = np.arange(5000, dtype=np.uint64) * 7919
    for n in (1, 2, 3, 8):
        want = np.array([lib.apm_owner_of_key(int(k), n) for k in keys], dtype=np.uint32)
        assert np.array_equal(sharding.owner_of_keys(keys, n), want)


def test_partition_is_a_disjoint_cover_in_arrival_order():
    svc, jvm, ts, el = _stream()
    parts = sharding.partition_columns(svc, jvm, ts, el, CFG["S"], 4)
    assert sum(p[0].size for p in parts) == svc.size
    seen = set()
    for p in parts:
        keys = set((p[1].astype(np.int64) * CFG["S"] + p[0]).tolist())
        assert not (keys & seen)
        seen |= keys
        assert np.all(np.diff(np.flatnonzero(np.isin(ts, p[2]))) >= 0)


def _run_rank(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    svc, jvm, ts, el = _stream()
    mine = sharding.partition_columns(svc, jvm, ts, el, CFG["S"], world)[rank]
    # ticks are global: replay the whole stream's tick positions, feeding only this rank's keys
    keep = np.zeros(CFG["S"] * CFG["J"], np.uint8)
    keys = np.arange(CFG["S"] * CFG["J"], dtype=np.uint64)
    keep[sharding.owner_of_keys(keys, world) == rank] = 1
    p = _pipeline()
    p.set_filter(keep)
    p.consume(svc, jvm, ts, el)
    fs = p.fs()
    alerts = fs[np.abs(fs["signal"]).sum(axis=1) > 0]
    assert mine[0].size == p.ingested
    gathered = [None] * world
    dist.all_gather_object(gathered, alerts.tobytes())
    merged = sharding.merge_alerts([np.frombuffer(b, dtype=ocapi.FS_DTYPE) for b in gathered])
    if rank == 0:
        q.put(merged.tobytes())
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gathered_alerts_equal_the_single_rank_run():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_run_rank, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = np.frombuffer(q.get(timeout=120), dtype=ocapi.FS_DTYPE)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = _pipeline()
    svc, jvm, ts, el = _stream()
    full.consume(svc, jvm, ts, el)
    fs = full.fs()
    want = fs[np.abs(fs["signal"]).sum(axis=1) > 0]
    assert want.size > 0 and got.size == want.size
    key = lambda a: a[np.lexsort((a["key"], a["lag"], a["tick_ts"]))]  # noqa: E731
    assert key(got).tobytes() == key(want).tobytes()


def test_global_tick_labels():
    svc, jvm, ts, el = _stream()
    labs = sharding.tick_labels(ts, CFG["bucket_ms"])
    p = _pipeline()
    p.consume(svc, jvm, ts, el)
    assert labs.size == p.ticks and labs[-1] == p.latest


def test_count_aware_placement_evens_out_a_two_tier_skew():
    """BASELINE config 5 (top 10 of 10 000 services = 80 % of the traffic) on 8 ranks: the hash puts up to 18.6 % of the
    transactions on one rank, the count-aware table 16 % - two hot services, the least a rank holding two of ten can have."""
    K = 10000
    counts = np.full(K, 0.2e9 / (K - 10))
    counts[:10] = 0.08e9
    for n in (2, 4, 8):
        hashed = np.bincount(sharding.owner_of_keys(np.arange(K), n), weights=counts, minlength=n)
        table = sharding.balanced_owner_table(counts, n)
        assert table.shape == (K,) and table.max() < n
        load = np.bincount(table, weights=counts, minlength=n)
        assert load.max() <= hashed.max()
        assert load.max() <= max(counts.sum() / n, 2 * counts[0]) * 1.001
    # with a per-key cost in the balance the ranks that hold two hot services keep fewer of the light keys
    t0 = sharding.balanced_owner_table(counts, 8)
    t1 = sharding.balanced_owner_table(counts, 8, key_cost=160e3)
    two_hot = np.flatnonzero(np.bincount(t1[:10], minlength=8) == 2)
    keys0, keys1 = np.bincount(t0, minlength=8), np.bincount(t1, minlength=8)
    assert two_hot.size == 2 and np.all(keys1[two_hot] > keys0[two_hot]) and np.all(keys1[two_hot] < keys1.max())
    cost = np.bincount(t1, weights=counts + 160e3, minlength=8)
    assert cost.max() / cost.mean() < 1.01
    # light keys stay with their hash owner wherever that rank has room
    uniform = np.full(K, 100.0)
    t = sharding.balanced_owner_table(uniform, 4)
    assert np.mean(t == sharding.owner_of_keys(np.arange(K), 4)) > 0.95
    assert np.array_equal(sharding.balanced_owner_table(np.zeros(16), 4), sharding.owner_of_keys(np.arange(16), 4))


def test_partition_with_an_owner_table_is_still_a_disjoint_cover():
    svc, jvm, ts, el = _stream()
    K = CFG["S"] * CFG["J"]
    key = jvm.astype(np.int64) * CFG["S"] + svc
    table = sharding.balanced_owner_table(np.bincount(key, minlength=K), 3)
    parts = sharding.partition_columns(svc, jvm, ts, el, CFG["S"], 3, owner_table=table)
    assert sum(p[0].size for p in parts) == svc.size
    for r, p in enumerate(parts):
        assert np.all(table[p[1].astype(np.int64) * CFG["S"] + p[0]] == r)
    sizes = np.array([p[0].size for p in parts])
    hashed = np.array([p[0].size for p in sharding.partition_columns(svc, jvm, ts, el, CFG["S"], 3)])
    assert sizes.max() <= hashed.max()


def test_owner_table_of_the_c_abi_equals_the_host_helper():
    lib = _capi.lib()
    rng = np.random.default_rng(7)
    cases = [(np.r_[np.full(10, 80_000_000), np.full(9990, 20_020)].astype(np.uint64), 8, 160_000),
             (np.r_[np.full(10, 80_000_000), np.full(9990, 20_020)].astype(np.uint64), 8, 0),
             (rng.zipf(1.3, 5000).astype(np.uint64), 3, 0), (rng.integers(0, 1000, 777).astype(np.uint64), 5, 50),
             (np.zeros(16, np.uint64), 4, 0), (np.array([5], np.uint64), 2, 0)]
    for counts, n, cost in cases:
        out = np.empty(counts.size, np.uint32)
        assert lib.apm_owner_table_build(counts.ctypes.data, counts.size, n, cost, out.ctypes.data) == 0
        assert np.array_equal(out, sharding.balanced_owner_table(counts, n, key_cost=cost)), (n, cost)
    assert lib.apm_owner_table_build(None, 4, 2, 0, None) != 0


def _local_ids(table_or_none, K, nranks):
    """Dense id of every key on its owner, in key order."""
    own = sharding.owner_of_keys(np.arange(K), nranks) if table_or_none is None else np.asarray(table_or_none)
    local = np.zeros(K, np.uint32)
    for r in range(nranks):
        idx = np.flatnonzero(own == r)
        local[idx] = np.arange(idx.size, dtype=np.uint32)
    return own, local


@pytest.mark.parametrize("nranks", [1, 2, 3, 8])
def test_native_partitioner_equals_the_numpy_mirror(nranks):
    """apm_partition_host (host arithmetic of the C ABI) against sharding.partition_columns + sharding.tick_marks."""
    svc, jvm, ts, el = make_stream(41, 5000, CFG["S"], CFG["J"], CFG["bucket_ms"], 120, late_frac=0.05, gap_at=(40, 47))
    K = CFG["S"] * CFG["J"]
    key = jvm.astype(np.int64) * CFG["S"] + svc
    for table in (None, sharding.balanced_owner_table(np.bincount(key, minlength=K), nranks)):
        parts = sharding.partition_columns(svc, jvm, ts, el, CFG["S"], nranks, owner_table=table)
        marks = sharding.tick_marks(svc, jvm, ts, CFG["bucket_ms"], CFG["S"], nranks, latest=0, owner_table=table)
        got, latest = sharding.partition_native(svc, jvm, ts, el, CFG["bucket_ms"], CFG["S"], CFG["J"], nranks,
                                                owner_table=table)
        assert latest == int((ts // CFG["bucket_ms"]).max())
        for r in range(nranks):
            for a, b in zip(got[r][0], parts[r]):
                assert np.array_equal(a, b)
            assert got[r][1].tobytes() == marks[r].tobytes()
            # every rank ticks at every global label, also while its shard is silent
            assert np.array_equal(got[r][1]["label"], sharding.tick_labels(ts, CFG["bucket_ms"]))
            assert np.all(np.diff(got[r][1]["pos"].astype(np.int64)) >= 0) and got[r][1]["pos"].max() <= parts[r][0].size
        # dense per-owner ids: service = the key's rank among its owner's keys, jvm = 0
        own, local = _local_ids(table, K, nranks)
        got2, _ = sharding.partition_native(svc, jvm, ts, el, CFG["bucket_ms"], CFG["S"], CFG["J"], nranks,
                                            owner_table=table, local_id=local)
        for r in range(nranks):
            pk = parts[r][1].astype(np.int64) * CFG["S"] + parts[r][0]
            assert np.array_equal(got2[r][0][0], local[pk]) and not got2[r][0][1].any()
            assert np.array_equal(got2[r][0][2], parts[r][2]) and np.array_equal(got2[r][0][3], parts[r][3])


def test_native_partitioner_carries_latest_and_routes_invalid_rows_to_rank_0():
    svc, jvm, ts, el = make_stream(43, 2000, CFG["S"], CFG["J"], CFG["bucket_ms"], 60)
    svc = svc.copy(); ts = ts.copy()
    svc[100] = CFG["S"] + 5          # service id out of range
    ts[200] = -1                     # negative timestamp: never ticks
    half = 1000
    a, latest = sharding.partition_native(svc[:half], jvm[:half], ts[:half], el[:half], CFG["bucket_ms"], CFG["S"], CFG["J"], 4)
    b, latest2 = sharding.partition_native(svc[half:], jvm[half:], ts[half:], el[half:], CFG["bucket_ms"], CFG["S"], CFG["J"], 4,
                                           latest=latest)
    whole, latest3 = sharding.partition_native(svc, jvm, ts, el, CFG["bucket_ms"], CFG["S"], CFG["J"], 4)
    assert latest2 == latest3
    for r in range(4):
        for i in range(4):
            assert np.array_equal(np.concatenate([a[r][0][i], b[r][0][i]]), whole[r][0][i])
        # marks of the second call are relative to it and continue above the first call's latestBucket
        joined = np.concatenate([a[r][1]["label"], b[r][1]["label"]])
        assert np.array_equal(joined, whole[r][1]["label"])
        assert np.array_equal(np.concatenate([a[r][1]["pos"], b[r][1]["pos"] + a[r][0][0].size]), whole[r][1]["pos"])
    assert (CFG["S"] + 5) in whole[0][0][0]                    # no key, no owner: rank 0 counts it as invalid
    assert sum(int((-1 == whole[r][0][2]).sum()) for r in range(4)) == 1  # a bad timestamp stays with its key's owner
    for r in range(1, 4):
        assert (CFG["S"] + 5) not in whole[r][0][0]


def test_native_partitioner_reports_full_outputs():
    import ctypes as C
    svc, jvm, ts, el = make_stream(44, 100, 4, 1, 10000, 10)
    outs = (_capi.PartitionOut * 2)()
    bufs = [(np.empty(10, np.uint32), np.empty(10, np.uint32), np.empty(10, np.int64), np.empty(10, np.int32),
             np.zeros(64, _capi.MARK_DTYPE)) for _ in range(2)]
    for r, b in enumerate(bufs):
        outs[r].cols = _capi.TxColumns(*(x.ctypes.data for x in b[:4]))
        outs[r].cap, outs[r].marks, outs[r].marks_cap = 10, b[4].ctypes.data, 64
    cols = _capi.TxColumns(svc.ctypes.data, jvm.ctypes.data, ts.ctypes.data, el.ctypes.data)
    st = _capi.lib().apm_partition_host(C.byref(cols), svc.size, 10000, 0, 4, 1, 2, None, None, outs, None)
    assert st == _capi.E_CAPACITY and b"full" in _capi.lib().apm_last_error(None)


def test_capped_placement_balances_transactions_with_even_key_counts():
    """80 k keys of Zipf(1) traffic on 8 ranks: the hash gives one rank 1.55 x the mean; the capped placement evens the
    transactions out while no rank holds more than 1 % above its share of the keys.  C ABI == numpy mirror."""
    lib = _capi.lib()
    K, n = 80_000, 8
    counts = np.rint(800e6 / (np.arange(1, K + 1) * np.log(K))).astype(np.uint64)
    out = np.empty(K, np.uint32)
    assert lib.apm_owner_table_build_capped(counts.ctypes.data, K, n, 0.01, out.ctypes.data) == 0
    tx = np.bincount(out, weights=counts.astype(np.float64), minlength=n)
    keys = np.bincount(out, minlength=n)
    hashed = np.bincount(sharding.owner_of_keys(np.arange(K), n), weights=counts.astype(np.float64), minlength=n)
    assert hashed.max() / hashed.mean() > 1.4
    assert tx.max() / tx.mean() < 1.01 and keys.max() <= np.ceil(K / n * 1.01)
    small = counts[:3000]
    o2 = np.empty(small.size, np.uint32)
    assert lib.apm_owner_table_build_capped(small.ctypes.data, small.size, 3, 0.05, o2.ctypes.data) == 0
    assert np.array_equal(o2, sharding.capped_owner_table(small, 3, 0.05))
    # two-tier skew (config 5): two ranks must hold two hot services; the cap leaves them the lightest keys
    c5 = np.r_[np.full(10, 80_000_000), np.full(9990, 20_020)].astype(np.uint64)
    o5 = np.empty(c5.size, np.uint32)
    assert lib.apm_owner_table_build_capped(c5.ctypes.data, c5.size, 8, 0.01, o5.ctypes.data) == 0
    assert np.bincount(o5, minlength=8).max() <= 1263 and sorted(np.bincount(o5[:10], minlength=8).tolist()) == [1] * 6 + [2, 2]
