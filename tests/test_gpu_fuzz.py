"""Seeded sweep over the configuration space: random window / buffer / bucket width / lags / percentiles /
thresholds / influences, random chunking of the call, late and unsorted arrivals, gaps - every record of
every tick bit-exact against the oracle.  Complements the hand-picked cases of test_gpu_parity.py."""
import numpy as np
import pytest

from tests.gpu_util import assert_fs_equal, assert_stats_equal, run_both
from tests.streams import make_stream

pytestmark = pytest.mark.gpu


def _case(seed):
    r = np.random.default_rng(1000 + seed)
    window = int(r.choice([1, 2, 3, 5, 12, 30, 40]))
    buffer = int(r.integers(0, min(7, 63 - window)))
    nlags = int(r.integers(1, 4))
    lags = tuple(sorted(int(x) for x in r.choice([3, 5, 8, 13, 21, 34, 60, 90], size=nlags, replace=False)))
    npct = int(r.integers(1, 5))
    pcts = tuple(float(x) for x in sorted(r.choice([0.0, 10.0, 25.0, 50.0, 75.0, 90.0, 95.0, 99.0, 99.9, 100.0], size=npct, replace=False)))
    zpct = (int(r.integers(0, npct)), int(r.integers(0, npct)))
    span = int(r.integers(max(lags) + window + buffer + 10, max(lags) + window + buffer + 120))
    return dict(seed=seed, n=int(r.integers(500, 30000)), S=int(r.integers(1, 40)), J=int(r.integers(1, 4)),
                bucket_ms=int(r.choice([1000, 10000, 60000])), window=window, buffer=buffer, span=span, lags=lags,
                thr=tuple(float(x) for x in r.uniform(0.5, 4.0, nlags)), infl=tuple(float(x) for x in r.choice([0.0, 0.1, 0.5, 1.0], nlags)),
                pcts=pcts, zpct=zpct, gap=None if r.random() < 0.6 else (span // 3, span // 3 + int(r.integers(1, 30))),
                log_cap=1 << 16, max_batch=int(r.choice([1 << 9, 1 << 12, 1 << 16])), out_ticks=2048)


INCREMENTAL = 1 | 2 | 4 | 8 | 16 | 32  # the default flags + APM_F_INCREMENTAL


@pytest.mark.parametrize("mode", ["exact", "incremental"])
@pytest.mark.parametrize("seed", range(32))
def test_random_configuration_matches_oracle(seed, mode):
    c = _case(seed)
    r = np.random.default_rng(seed)
    cols = make_stream(c["seed"], c["n"], c["S"], c["J"], c["bucket_ms"], c["span"], gap_at=c["gap"],
                       shift_services=(0, 2), late_frac=float(r.choice([0.0, 0.01, 0.1])),
                       late_ms=int(c["bucket_ms"] * r.integers(1, 12)), sorted_time=bool(r.random() < 0.8),
                       elapsed_hi=int(r.choice([300, 60000, 2_000_000])))
    n = cols[0].size
    chunks = sorted(int(x) for x in r.integers(0, n, size=int(r.integers(0, 6)))) or None
    kw = dict(flags=INCREMENTAL) if mode == "incremental" else {}
    eng, orc, st, fs = run_both(c, cols, chunks=chunks, **kw)
    assert orc.ticks > 0
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs(), exact=mode == "exact")  # incremental: signals / values exact, mean and bounds to 1e-9
    cnt = eng.counters()
    assert cnt.ticks == orc.ticks
    eng.close()
