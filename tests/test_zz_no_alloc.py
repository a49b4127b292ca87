"""apm_sorted_percentiles_noalloc - the percentile operator for callers that cannot allocate
device memory at call time (GPU memory fully pre-reserved by the host process, handlers on
a latency budget, deployments that forbid allocation after start).  Its scratch is a fixed
array inside the library,
shared by all calls and never cleared between them, so the test calls it twice in a row -
a large input first, then a smaller one that leaves stale words behind it in the scratch -
and requires the same answers as the allocating entry point, bit for bit (integer sort,
percentiles that are integers or x.5).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from apmbackend_b200 import _capi, array_ops

pytestmark = pytest.mark.gpu

PCTS = (75.0, 95.0, 50.0, 99.0)


def _values(seed, n):
    rng = np.random.default_rng(seed)
    v = np.clip(np.rint(np.exp(rng.normal(5.0, 0.9, n))), 0, 900000).astype(np.int32)
    v[: n // 4] = v[0] if n else 0
    return v


def test_noalloc_matches_allocating_entry_point_on_two_calls_in_a_row():
    for seed, n in ((1, 200_000), (2, 4097)):  # second call reuses scratch the first one dirtied
        v = _values(seed, n)
        want_pct, want_sorted = array_ops.sorted_percentiles(v, PCTS)
        got_pct, got_sorted = array_ops.sorted_percentiles_noalloc(v, PCTS)
        assert np.array_equal(got_sorted, want_sorted)
        assert np.array_equal(got_pct.view(np.uint64), want_pct.view(np.uint64))
        assert np.array_equal(got_sorted, np.sort(v))


def test_noalloc_small_and_empty_inputs():
    for n in (0, 1, 5, 20):
        v = _values(n + 10, n)
        want_pct, want_sorted = array_ops.sorted_percentiles(v, PCTS)
        got_pct, got_sorted = array_ops.sorted_percentiles_noalloc(v, PCTS)
        assert np.array_equal(got_sorted, want_sorted)
        assert np.array_equal(np.isnan(got_pct), np.isnan(want_pct))
        assert np.array_equal(got_pct[~np.isnan(got_pct)], want_pct[~np.isnan(want_pct)])


def test_noalloc_rejects_sizes_beyond_its_maximum():
    lib = _capi.lib()
    out = (C.c_double * 1)()
    p = (C.c_double * 1)(75.0)
    fake = C.c_void_p(256)  # never dereferenced: the size check comes first
    st = lib.apm_sorted_percentiles_noalloc(0, fake, array_ops.NOALLOC_MAX_VALUES + 1, p, 1, out, None)
    assert st == _capi.E_CAPACITY
    assert b"APM_NOALLOC_MAX_VALUES" in lib.apm_last_error(None)
    v = _values(3, array_ops.NOALLOC_MAX_VALUES)  # exactly the maximum is accepted
    got_pct, got_sorted = array_ops.sorted_percentiles_noalloc(v, PCTS)
    assert np.array_equal(got_sorted, np.sort(v))


def test_noalloc_call_leaves_the_free_device_memory_untouched():
    """What the entry point promises: once the library's kernels and its static scratch are on the device
    (apm_preload_kernels, or simply an earlier call), a call asks the CUDA allocator for nothing - the free memory
    cudaMemGetInfo reports is the same before and after, also for the second call in a row.  The allocating entry
    point is measured the same way, to show that the probe would see an allocation that stayed."""
    from apmbackend_b200.synthetic import DeviceBuffer
    lib, S = _capi.lib(), _capi.synth()
    assert lib.apm_preload_kernels(0) == _capi.OK
    v = _values(7, 150_000)
    d_in, d_out = DeviceBuffer(v.nbytes).upload(v), DeviceBuffer(v.nbytes)
    pcts = (C.c_double * 4)(*PCTS)
    out = (C.c_double * 4)()

    def free_now():
        S.synth_sync()
        f, t = C.c_uint64(0), C.c_uint64(0)
        assert S.synth_mem_info(C.byref(f), C.byref(t)) == 0
        return f.value

    assert lib.apm_sorted_percentiles_noalloc(0, C.c_void_p(d_in.ptr), v.size, pcts, 4, out, C.c_void_p(d_out.ptr)) == _capi.OK  # warm
    before = free_now()
    for _ in range(2):
        assert lib.apm_sorted_percentiles_noalloc(0, C.c_void_p(d_in.ptr), v.size, pcts, 4, out, C.c_void_p(d_out.ptr)) == _capi.OK
        assert free_now() == before
    assert np.array_equal(d_out.download(np.int32, v.size), np.sort(v))
    d_in.free(); d_out.free()
