"""Host-side key sharding for the multi-GPU path: every (server, service) key belongs to
exactly one rank, `owner = splitmix64(key) mod nranks` (the same function as
apm_owner_of_key in the C ABI), and the host bins transactions by owner at ingest.  No
statistic has a cross-key term (stream_calc_stats.js:159-202, stream_calc_z_score.js:197-308),
so the ranks never exchange data-path state; only their compact alert records are gathered."""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def owner_of_keys(keys, nranks: int) -> np.ndarray:
    """Vectorised apm_owner_of_key."""
    k = np.asarray(keys, dtype=np.uint64).copy()
    with np.errstate(over="ignore"):
        k += np.uint64(0x9E3779B97F4A7C15)
        k = (k ^ (k >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        k = (k ^ (k >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        k ^= k >> np.uint64(31)
    return (k % np.uint64(max(nranks, 1))).astype(np.uint32)


def balanced_owner_table(key_counts, nranks: int, key_cost: float = 0.0) -> np.ndarray:
    """Count-aware placement for skewed traffic (BASELINE config 5: ten services carry 80 % of the transactions):
    owner[key] for every dense key id, from the transactions each key had in an earlier interval.  Keys that alone
    exceed 1 / (8 nranks) of the traffic are placed longest-first on the least loaded rank; the many light keys keep
    their hash owner unless that rank is already above the mean, in which case they go to the least loaded one.
    A key is never split, so a rank cannot get below its heaviest key's share; rings do not migrate - the table is
    meant to be chosen when a deployment starts (or from a resume file), not changed under load.
    key_cost: what holding one key costs per interval, in transactions (the exact-order baseline walks every key's
    lag lists each tick whatever its traffic): the load that is balanced is count + key_cost per key."""
    counts = np.asarray(key_counts, dtype=np.float64) + float(key_cost)
    n = max(int(nranks), 1)
    owner = owner_of_keys(np.arange(counts.size, dtype=np.uint64), n)
    total = float(counts.sum())
    if n == 1 or total <= 0:
        return owner
    load = np.zeros(n)
    heavy = np.flatnonzero(counts > total / (8.0 * n))
    for k in heavy[np.argsort(-counts[heavy], kind="stable")]:
        r = int(np.argmin(load))
        owner[k] = r
        load[r] += counts[k]
    light = np.ones(counts.size, bool)
    light[heavy] = False
    mean = total / n
    for k in np.flatnonzero(light)[np.argsort(-counts[light], kind="stable")]:
        r = int(owner[k])
        if load[r] + counts[k] > mean:
            r = int(np.argmin(load))
            owner[k] = r
        load[r] += counts[k]
    return owner


def capped_owner_table(key_counts, nranks: int, key_slack: float = 0.01) -> np.ndarray:
    """apm_owner_table_build_capped: transactions balanced longest-first, every rank's key count within (1 + key_slack) of
    the mean (its ring memory and baseline walk follow the keys, its ingest / sort / PCIe time the transactions)."""
    import math
    counts = np.asarray(key_counts, dtype=np.uint64)
    n = max(int(nranks), 1)
    cap = math.ceil(counts.size / n * (1.0 + max(float(key_slack), 0.0)))
    owner = np.zeros(counts.size, np.uint32)
    load = [0] * n
    held = [0] * n
    for k in np.argsort(-counts.astype(np.int64), kind="stable"):
        best = -1
        for r in range(n):
            if held[r] < cap and (best < 0 or load[r] < load[best]):
                best = r
        owner[k] = best
        load[best] += int(counts[k])
        held[best] += 1
    return owner


def partition_columns(service_id, jvm_id, end_ts_ms, elapsed_ms, num_services: int, nranks: int, owner_table=None):
    """Splits one arrival-ordered stream into per-rank streams (arrival order kept inside each).
    Returns a list of (service_id, jvm_id, end_ts_ms, elapsed_ms) per rank.  owner_table (balanced_owner_table)
    replaces the hash owner when given."""
    svc = np.asarray(service_id, dtype=np.uint32)
    jvm = np.asarray(jvm_id, dtype=np.uint32)
    key = jvm.astype(np.uint64) * np.uint64(num_services) + svc.astype(np.uint64)
    own = owner_of_keys(key, nranks) if owner_table is None else np.asarray(owner_table, dtype=np.uint32)[key]
    out = []
    for r in range(nranks):
        m = own == r
        out.append((svc[m], jvm[m], np.asarray(end_ts_ms)[m], np.asarray(elapsed_ms)[m]))
    return out


def tick_labels(end_ts_ms, bucket_ms: int, latest: int = 0) -> np.ndarray:
    """The bucket labels at which the GLOBAL stream ticks (running maximum increases).  Ticks are
    global: a rank whose shard is silent in a bucket still has to tick, so the host broadcasts
    these labels and every rank calls apm_tick for each of them at the same stream position."""
    lab = np.asarray(end_ts_ms, dtype=np.int64) // bucket_ms
    run = np.maximum.accumulate(np.concatenate([[latest], lab]))
    return np.unique(run[1:][run[1:] > run[:-1]])


def tick_marks(service_id, jvm_id, end_ts_ms, bucket_ms: int, num_services: int, nranks: int, latest: int = 0,
               owner_table=None):
    """Per rank, where the GLOBAL ticks fall inside that rank's sub-stream: a structured (pos, label) array per rank for
    Engine.set_tick_marks.  pos = the rank's transactions that precede the trigger (numpy mirror of apm_partition_host;
    assumes every transaction passes the apm_tx_columns preconditions)."""
    from ._capi import MARK_DTYPE
    ts = np.asarray(end_ts_ms, dtype=np.int64)
    lab = np.where(ts >= 0, ts // bucket_ms, -1)
    run = np.maximum.accumulate(np.concatenate([[latest], lab]))
    trig = np.flatnonzero(run[1:] > run[:-1])          # global positions of the triggers
    key = np.asarray(jvm_id, dtype=np.uint64) * np.uint64(num_services) + np.asarray(service_id, dtype=np.uint64)
    own = owner_of_keys(key, nranks) if owner_table is None else np.asarray(owner_table, dtype=np.uint32)[key]
    out = []
    for r in range(nranks):
        before = np.concatenate([[0], np.cumsum(own == r)])  # before[i] = rank-r transactions among the first i
        m = np.zeros(trig.size, dtype=MARK_DTYPE)
        m["pos"] = before[trig]
        m["label"] = lab[trig]
        out.append(m)
    return out


def partition_native(service_id, jvm_id, end_ts_ms, elapsed_ms, bucket_ms: int, num_services: int, num_jvms: int,
                     nranks: int, latest: int = 0, owner_table=None, local_id=None):
    """apm_partition_host through the C ABI: [(columns, marks)] per rank and latestBucket after the stream."""
    import ctypes as C
    from . import _capi
    svc = np.ascontiguousarray(service_id, dtype=np.uint32)
    jvm = np.ascontiguousarray(jvm_id, dtype=np.uint32)
    ts = np.ascontiguousarray(end_ts_ms, dtype=np.int64)
    el = np.ascontiguousarray(elapsed_ms, dtype=np.int32)
    n = svc.size
    lab = ts // bucket_ms
    marks_cap = int(np.unique(lab).size) + 1
    outs = (_capi.PartitionOut * nranks)()
    bufs = []
    for r in range(nranks):
        b = (np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(n, np.int64), np.empty(n, np.int32),
             np.zeros(marks_cap, _capi.MARK_DTYPE))
        bufs.append(b)
        outs[r].cols = _capi.TxColumns(b[0].ctypes.data, b[1].ctypes.data, b[2].ctypes.data, b[3].ctypes.data)
        outs[r].cap = n
        outs[r].marks = b[4].ctypes.data
        outs[r].marks_cap = marks_cap
    ot = None if owner_table is None else np.ascontiguousarray(owner_table, dtype=np.uint32)
    li = None if local_id is None else np.ascontiguousarray(local_id, dtype=np.uint32)
    cols = _capi.TxColumns(svc.ctypes.data, jvm.ctypes.data, ts.ctypes.data, el.ctypes.data)
    latest_out = C.c_int64(0)
    st = _capi.lib().apm_partition_host(C.byref(cols), n, int(bucket_ms), int(latest), int(num_services), int(num_jvms),
                                        int(nranks), ot.ctypes.data if ot is not None else None,
                                        li.ctypes.data if li is not None else None, outs, C.byref(latest_out))
    if st != _capi.OK:
        raise _capi.ApmError(st, (_capi.lib().apm_last_error(None) or b"").decode())
    res = []
    for r in range(nranks):
        k, m = int(outs[r].n), int(outs[r].num_marks)
        b = bufs[r]
        res.append(((b[0][:k].copy(), b[1][:k].copy(), b[2][:k].copy(), b[3][:k].copy()), b[4][:m].copy()))
    return res, int(latest_out.value)


def merge_alerts(per_rank_alerts):
    """What the all-gather delivers: rank-major concatenation, (tick, lag, key)-sorted inside a rank."""
    parts = []
    for a in per_rank_alerts:
        order = np.lexsort((a["key"], a["lag"], a["tick_ts"]))
        parts.append(a[order])
    return np.concatenate(parts) if parts else np.zeros(0)
