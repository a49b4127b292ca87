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


def partition_columns(service_id, jvm_id, end_ts_ms, elapsed_ms, num_services: int, nranks: int):
    """Splits one arrival-ordered stream into per-rank streams (arrival order kept inside each).
    Returns a list of (service_id, jvm_id, end_ts_ms, elapsed_ms) per rank."""
    svc = np.asarray(service_id, dtype=np.uint32)
    jvm = np.asarray(jvm_id, dtype=np.uint32)
    key = jvm.astype(np.uint64) * np.uint64(num_services) + svc.astype(np.uint64)
    own = owner_of_keys(key, nranks)
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


def merge_alerts(per_rank_alerts):
    """What the all-gather delivers: rank-major concatenation, (tick, lag, key)-sorted inside a rank."""
    parts = []
    for a in per_rank_alerts:
        order = np.lexsort((a["key"], a["lag"], a["tick_ts"]))
        parts.append(a[order])
    return np.concatenate(parts) if parts else np.zeros(0)
