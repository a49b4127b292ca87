"""One host process, one context per GPU (the reference stage is one Node process,
apm_manager.js:329-347; the drop-in drives one pinned ring per GPU from it).  Every entry
point makes its context's device current, and the kernels' attributes are set on each
device, so contexts on devices 0 and 1 can be used in any interleaving.  Skipped on a
box with fewer than two GPUs."""
import numpy as np
import pytest

from apmbackend_b200 import Engine, _capi
from oracle import capi as ocapi
from tests.gpu_util import assert_fs_equal, assert_stats_equal, engine_config, oracle_config
from tests.streams import make_stream
from tests.test_oracle_twins import CASES

pytestmark = pytest.mark.gpu


def _need_two():
    if _capi.device_count() < 2:
        pytest.skip("needs two CUDA devices")


def _oracle(c, cols):
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    orc.consume(*cols)
    return orc


def test_two_contexts_on_two_devices_interleaved():
    _need_two()
    ca = dict(CASES[0])
    cb = dict(CASES[1])
    cols_a = make_stream(ca["seed"], ca["n"], ca["S"], ca["J"], ca["bucket_ms"], ca["span"], shift_services=(0, 2))
    cols_b = make_stream(cb["seed"] + 50, cb["n"], cb["S"], cb["J"], cb["bucket_ms"], cb["span"], shift_services=(1,))
    ea = Engine(engine_config(ca, device=0))
    eb = Engine(engine_config(cb, device=1))  # the second device sets its own kernel attributes
    st_a, fs_a, st_b, fs_b = [], [], [], []
    na, nb = cols_a[0].size, cols_b[0].size
    cuts = 5
    for i in range(cuts):
        sa = slice(i * na // cuts, (i + 1) * na // cuts)
        sb = slice(i * nb // cuts, (i + 1) * nb // cuts)
        ea.process_host(*(x[sa] for x in cols_a))
        eb.process_host(*(x[sb] for x in cols_b))
        st_b.append(eb.fetch_stats().copy()); fs_b.append(eb.fetch_fs().copy())  # fetch in the other order
        st_a.append(ea.fetch_stats().copy()); fs_a.append(ea.fetch_fs().copy())
        eb.release_ticks(); ea.release_ticks()
    oa, ob = _oracle(ca, cols_a), _oracle(cb, cols_b)
    assert_stats_equal(np.concatenate(st_a), oa.stats()); assert_fs_equal(np.concatenate(fs_a), oa.fs())
    assert_stats_equal(np.concatenate(st_b), ob.stats()); assert_fs_equal(np.concatenate(fs_b), ob.fs())
    assert ea.counters().ticks == oa.ticks and eb.counters().ticks == ob.ticks
    ea.close(); eb.close()


def test_stats_only_table_kernel_on_the_second_device():
    """reduce_kernel_v3<true> needs ~98 KB of dynamic shared memory: the limit must have been raised on device 1 too."""
    _need_two()
    c = dict(CASES[0], pcts=(), lags=(), thr=(), infl=())
    cols = make_stream(5, 20000, c["S"], c["J"], c["bucket_ms"], c["span"])
    res = []
    for dev in (0, 1):
        e = Engine(engine_config(c, device=dev, flags=_capi.F_KEEP_STATS))
        e.process_host(*cols)
        res.append(e.fetch_stats().copy())
        e.close()
    assert res[0].size > 0 and res[0].tobytes() == res[1].tobytes()
