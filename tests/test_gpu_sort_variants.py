"""Every form of the bucket sort behind the percentiles gives the same records: the radix passes (the default) and the
planned partition + group sort (APM_SORT_GROUP=1) - packed 32-bit words (key + value fit
32 bits) and 64-bit words, the compile-time 8-bit digit and run-time digit widths, the three-kernel fallback.
The in-process run (default settings) is checked against the oracle; each variant then runs in a child process -
the sort reads its environment knobs once per process - and must reproduce that run's records bit for bit."""
import hashlib
import os
import subprocess
import sys

import numpy as np
import pytest

from tests.gpu_util import assert_fs_equal, assert_stats_equal, run_both, sort_records
from tests.streams import make_stream

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = {
    # 10 000 keys x durations < 2^17: 14 + 17 = 31 significant bits -> packed u32 words, 4 x 8 bits (3 x 11 on request)
    "narrow": dict(seed=31, n=70000, S=5000, J=2, bucket_ms=10000, window=2, buffer=0, span=8, lags=(3,), thr=(1.5,),
                   infl=(0.2,), pcts=(50.0, 95.0, 99.0), gap=None, log_cap=1 << 15, max_batch=1 << 17, elapsed_hi=100000),
    # 70 000 keys x durations up to 2^20: 17 + 20 = 37 bits -> u64 words, 5 passes
    "wide": dict(seed=32, n=70000, S=70000, J=1, bucket_ms=10000, window=2, buffer=0, span=8, lags=(3,), thr=(1.5,),
                 infl=(0.2,), pcts=(50.0, 95.0, 99.0), gap=None, log_cap=1 << 15, max_batch=1 << 17, elapsed_hi=1 << 20),
}
# APM_SORT_GROUP=1: every bucket after a context's first is planned from the bucket before it (partition + shared-memory
# group sorts, k3_groupsort.cu) instead of going down the radix passes the other knobs select
VARIANTS = [{"APM_SORT_GROUP": "1"}, {"APM_SORT_NARROW": "0"}, {"APM_SORT_BITS": "11"}, {"APM_SORT_BITS": "10", "APM_SORT_NARROW": "0"},
            {"APM_SORT_LEGACY": "1"}]


def _stream(c):
    svc, jvm, ts, el = make_stream(c["seed"], c["n"], c["S"], c["J"], c["bucket_ms"], c["span"], shift_services=(0, 2),
                                   late_frac=0.02, late_ms=15000, elapsed_hi=c["elapsed_hi"])
    el[::997] = c["elapsed_hi"] - 1 - np.arange(el[::997].size, dtype=np.int32)  # the value field really is that wide
    return svc, jvm, ts, el


def _digest(st, fs):
    h = hashlib.sha256()
    h.update(sort_records(st).tobytes())
    h.update(sort_records(fs, True).tobytes())
    return h.hexdigest()


def _child(name):
    """python -m tests.test_gpu_sort_variants <case>: engine only, prints the digest of its records."""
    from apmbackend_b200 import Engine
    from tests.gpu_util import engine_config
    c = CASES[name]
    svc, jvm, ts, el = _stream(c)
    eng = Engine(engine_config(c))
    eng.process_host(svc, jvm, ts, el)
    print("DIGEST", _digest(eng.fetch_stats().copy(), eng.fetch_fs().copy()))
    eng.close()


@pytest.mark.parametrize("name", sorted(CASES))
def test_sort_variants_give_identical_records(name):
    c = CASES[name]
    cols = _stream(c)
    per_bucket = np.unique(cols[2] // c["bucket_ms"], return_counts=True)[1]
    assert (per_bucket > 4096).any(), "buckets must be above the one-CTA sort's limit"
    eng, orc, st, fs = run_both(c, cols)
    assert_stats_equal(st, orc.stats())
    assert_fs_equal(fs, orc.fs())
    eng.close()
    want = _digest(st, fs)
    for env in VARIANTS:
        e = dict(os.environ, **env)
        out = subprocess.run([sys.executable, "-m", "tests.test_gpu_sort_variants", name], cwd=ROOT, env=e, capture_output=True,
                             text=True, timeout=300)
        assert out.returncode == 0, (env, out.stderr[-2000:])
        got = [ln.split()[1] for ln in out.stdout.splitlines() if ln.startswith("DIGEST")]
        assert got == [want], (env, got, want)


if __name__ == "__main__":
    _child(sys.argv[1])
