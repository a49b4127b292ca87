"""The reference's Array.prototype helpers (util_methods.js:430-446) as device operators.
`binaryConcat` + `calcPercentile` become one call: sort the durations on the GPU, then apply
the reference's index rule (util_methods.js:112-142)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from .synthetic import DeviceBuffer


def _call(fn, device, values, percentiles, want_sorted=True):
    v = np.ascontiguousarray(values, dtype=np.int32)
    p = np.ascontiguousarray(percentiles, dtype=np.float64)
    out = np.full(p.size, np.nan)
    d_in = DeviceBuffer(max(v.nbytes, 4)).upload(v) if v.size else DeviceBuffer(4)
    d_out = DeviceBuffer(max(v.nbytes, 4)) if want_sorted else None
    st = fn(device, C.c_void_p(d_in.ptr), v.size, p.ctypes.data, p.size, out.ctypes.data,
            C.c_void_p(d_out.ptr) if d_out else None)
    if st != _capi.OK:
        raise _capi.ApmError(st, (_capi.lib().apm_last_error(None) or b"").decode())
    srt = d_out.download(np.int32, v.size) if d_out and v.size else np.zeros(0, np.int32)
    d_in.free()
    if d_out:
        d_out.free()
    return out, srt


def sorted_percentiles(values, percentiles=(75.0, 95.0), device: int = 0):
    """Returns (percentile values, ascending copy of `values`)."""
    return _call(_capi.lib().apm_sorted_percentiles, device, values, percentiles)


NOALLOC_MAX_VALUES = 262144


def sorted_percentiles_noalloc(values, percentiles=(75.0, 95.0), device: int = 0):
    """Same result as sorted_percentiles, for hosts that cannot allocate device memory when
    they need it (all GPU memory already handed to the process's own pool, no room in the
    latency budget for cudaMalloc's synchronisation, a no-allocation-after-start policy): the
    operator's scratch is a fixed
    array inside the library.  At most NOALLOC_MAX_VALUES values; calls must not overlap on
    one GPU.  (This Python helper still stages `values` in device buffers of its own - only
    the library call is allocation-free.)"""
    return _call(_capi.lib().apm_sorted_percentiles_noalloc, device, values, percentiles)
