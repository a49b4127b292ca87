"""The alert all-gather over NCCL with real ranks: two contexts on two GPUs of one process (one host thread per rank,
as NCCL wants for communicators that share a process), each fed its key shard and the global tick marks.  What every
rank receives - on demand and in the streamed per-tick mode - must be the alert set of one unsharded context.
Skipped with fewer than two GPUs; bench.py's sharded_parity leg repeats it with one process per rank."""
import threading

import numpy as np
import pytest

from apmbackend_b200 import Engine, _capi, sharding
from tests.gpu_util import engine_config
from tests.streams import make_stream

pytestmark = pytest.mark.gpu

C0 = dict(S=24, J=2, bucket_ms=10000, window=3, buffer=1, lags=(6, 15), thr=(1.5, 2.5), infl=(0.3, 0.0), pcts=(75.0, 95.0))


def _rank_thread(r, world, uid, c, part, local_count, mode, calls, out, errs):
    try:
        cc = dict(c, S=max(local_count, 1), J=1)
        eng = Engine(engine_config(cc, device=r))
        eng.comm_init(r, world, uid)
        got = []
        if mode == "bound":
            eng.bind_gather_output(1 << 16)
        elif mode == "small_block":
            eng.comm_set_capacity(4)  # far too small: every rank must see the same error
        (svc, jvm, ts, el), marks = part
        n = svc.size
        bounds = np.linspace(0, n, calls + 1).astype(int)
        for a, b in zip(bounds[:-1], bounds[1:]):
            m = marks[(marks["pos"] >= a) & ((marks["pos"] < b) | (b == n))].copy()
            m["pos"] -= np.uint64(a)
            eng.set_tick_marks(m)
            eng.process_host(svc[a:b], jvm[a:b], ts[a:b], el[a:b])
            try:
                got.append(eng.gather_alerts(1 << 16, pinned=(mode != "pageable")).copy())
            except _capi.ApmError as e:
                got.append(e.status)
            eng.release_ticks()
        out[r] = got
        eng.close()
    except Exception as e:  # noqa: BLE001
        errs.append((r, repr(e)))


def _run(mode, calls=3, world=2):
    if _capi.device_count() < world:
        pytest.skip("needs two CUDA devices")
    c = dict(C0)
    cols = make_stream(71, 8000, c["S"], c["J"], c["bucket_ms"], 120, late_frac=0.05, shift_services=(0, 1, 2, 3, 4, 7))
    K = c["S"] * c["J"]
    own = sharding.owner_of_keys(np.arange(K), world)
    local = np.zeros(K, np.uint32)
    glob = []
    for r in range(world):
        idx = np.flatnonzero(own == r)
        local[idx] = np.arange(idx.size)
        glob.append(idx)
    parts, _ = sharding.partition_native(*cols, c["bucket_ms"], c["S"], c["J"], world, local_id=local)
    uid = Engine.comm_unique_id()
    out, errs = [None] * world, []
    th = [threading.Thread(target=_rank_thread, args=(r, world, uid, c, parts[r], glob[r].size, mode, calls, out, errs))
          for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    assert not errs, errs
    return c, cols, glob, out


def _unsharded_alerts(c, cols):
    eng = Engine(engine_config(c))
    eng.process_host(*cols)
    al = eng.fetch_alerts().copy()
    eng.close()
    return al


def _globalise(recs, glob):
    recs = recs.copy()
    for r, g in enumerate(glob):
        m = recs["pad"][:, 2] == r
        recs["key"][m] = g[recs["key"][m]]
    recs["pad"][:, 2] = 0
    return recs[np.lexsort((recs["key"], recs["lag"], recs["tick_ts"]))]


@pytest.mark.parametrize("mode", ["pinned", "pageable", "bound"])
def test_gathered_alerts_equal_the_unsharded_run(mode):
    c, cols, glob, out = _run(mode)
    want = _unsharded_alerts(c, cols)
    want = want[np.lexsort((want["key"], want["lag"], want["tick_ts"]))]
    assert want.size > 0
    for r in range(2):
        got = _globalise(np.concatenate(out[r]), glob)
        assert got.tobytes() == want.tobytes(), (mode, r, got.size, want.size)


def test_an_overflowing_block_is_an_error_on_every_rank_not_a_hang():
    _, _, _, out = _run("small_block", calls=1)
    assert out[0] == [_capi.E_CAPACITY] and out[1] == [_capi.E_CAPACITY]
