"""The walk for a few hundred lists (baseline_few_kernel, k4_baseline.cu): every lag of a tick in ONE launch, per 32
columns a loader warp and two adder warps that take the 96-slot chunks in turn and hand the running sum over through
shared memory.  It is what the reference's own deployment shape runs (400 keys, lags 360 + 8640), so its records must
equal the oracle's bit for bit - lists that are filling, full, wrapping around the ring (head in the middle of a
chunk), lags shorter than one chunk and of exactly one or two chunks, lists with undefined entries - and
equal those of the per-lag kernels it replaces (APM_K4_FEW=0)."""
import os

import numpy as np
import pytest

from apmbackend_b200 import Engine
from tests.gpu_util import assert_fs_equal, assert_stats_equal, engine_config, oracle_config
from tests.streams import make_stream
from oracle import capi as ocapi

pytestmark = pytest.mark.gpu


def _engine(c, few):
    old = os.environ.get("APM_K4_FEW")
    os.environ["APM_K4_FEW"] = "1" if few else "0"  # read when the context is created
    try:
        return Engine(engine_config(c))
    finally:
        if old is None:
            os.environ.pop("APM_K4_FEW", None)
        else:
            os.environ["APM_K4_FEW"] = old


def _run(eng, cols, calls):
    n = cols[0].size
    st, fs, al = [], [], []
    bounds = np.linspace(0, n, calls + 1).astype(int)
    for a, b in zip(bounds[:-1], bounds[1:]):
        eng.process_host(*(x[a:b] for x in cols))
        st.append(eng.fetch_stats().copy())
        fs.append(eng.fetch_fs().copy())
        al.append(eng.fetch_alerts().copy())
        eng.release_ticks()
    cnt = eng.counters()
    eng.close()
    return np.concatenate(st), np.concatenate(fs), np.concatenate(al), cnt.kernel_launches, cnt.ticks


@pytest.mark.parametrize("S,J,lags,span,calls", [
    (40, 2, (5, 130), 420, 3),         # a lag far below one chunk, one of 1.4 chunks; the ring wraps three times
    (200, 2, (96, 192, 333), 700, 2),  # lags of exactly one and two chunks and a ragged one; three lags in one launch
    (390, 1, (7, 61, 200, 401), 900, 4),  # four lags x 37 column groups = 148 CTAs: one CTA per SM
    (1000, 1, (50, 300), 640, 2),      # two lags x 94 column groups = 188 CTAs: more than the SMs, a second wave
], ids=["two-lags", "chunk-multiples", "four-lags-148-ctas", "two-waves"])
def test_few_list_walk_equals_the_oracle_and_the_per_lag_kernels(S, J, lags, span, calls):
    c = dict(S=S, J=J, bucket_ms=10000, window=3, buffer=1, pcts=(75.0, 95.0), lags=lags, thr=tuple(1.5 + 0.5 * i for i in range(len(lags))),
             infl=tuple(0.1 * i for i in range(len(lags))), log_cap=1 << 14, max_batch=1 << 15, out_ticks=span + 8)
    n = 40 * span
    svc, jvm, ts, el = make_stream(211, n, S, J, c["bucket_ms"], span, shift_services=(0, 3), late_frac=0.02, late_ms=15000)
    # two services fall silent for a stretch: their ticks push undefined entries, and the chunks that hold those take
    # the walk's entry-by-entry path (average() skips them)
    lab = ts // c["bucket_ms"]
    lo, hi = lab.min() + span // 3, lab.min() + span // 3 + 40
    quiet = (lab >= lo) & (lab < hi) & (svc < 2)
    svc = svc.copy()
    svc[quiet] = 5
    cols = (svc, jvm, ts, el)
    st, fs, al, launches, ticks = _run(_engine(c, True), cols, calls)
    st0, fs0, al0, launches0, ticks0 = _run(_engine(c, False), cols, calls)
    assert st.tobytes() == st0.tobytes() and fs.tobytes() == fs0.tobytes()
    # the alert blocks, which the walk's last CTA compacts itself: the signalled records in record order
    assert al.size > 0 and al.tobytes() == al0.tobytes()
    sig = fs[(fs["signal"] != 0).any(axis=1)]
    assert al.size == sig.size and np.array_equal(np.sort(al, order=["tick_ts", "lag", "key"]), np.sort(sig, order=["tick_ts", "lag", "key"]))
    assert ticks == ticks0 > 0
    # one launch per tick instead of one per lag + one for the alert compaction
    assert launches0 - launches == ticks * len(lags), (launches0, launches, ticks)
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    orc.consume(*cols)
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())
    full = fs[(fs["lag"] == lags[-1]) & ~np.isnan(fs["avg"][:, 0])]
    assert full.size > 0 and (fs["signal"] != 0).any()  # the longest list did fill up, and something signalled
