"""Synthetic transaction streams for bench.py and the tests: Zipf service
popularity, lognormal-shaped durations, near-sorted event time.  The generator is
counter-based and integer-only (csrc/synth_common.h), so the device and the host twin give
identical columns.  Bench/test support - not part of the drop-in API."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _capi
from ._capi import TxColumns


class SynthParams(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n", C.c_uint64), ("first", C.c_uint64),
                ("num_services", C.c_uint32), ("num_jvms", C.c_uint32),
                ("t0_ms", C.c_int64), ("span_ms", C.c_int64),
                ("jitter_ms", C.c_uint32), ("elapsed_max", C.c_uint32),
                ("shift_mod", C.c_uint32), ("shift_mul", C.c_uint32),
                ("q_size", C.c_uint32), ("pad_", C.c_uint32),
                ("cdf", C.c_void_p), ("base", C.c_void_p), ("q", C.c_void_p)]


def _chk(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what}: CUDA error {rc}: {_capi.synth().synth_error_string(rc).decode()}")


class DeviceBuffer:
    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        _chk(_capi.synth().synth_malloc(C.byref(p), max(self.nbytes, 1)), "cudaMalloc")
        self.ptr = p.value

    def upload(self, arr: np.ndarray):
        a = np.ascontiguousarray(arr)
        assert a.nbytes <= self.nbytes
        _chk(_capi.synth().synth_h2d(self.ptr, a.ctypes.data, a.nbytes), "H2D")
        return self

    def download(self, dtype, count, offset_bytes=0) -> np.ndarray:
        out = np.empty(count, dtype=dtype)
        _chk(_capi.synth().synth_d2h(out.ctypes.data, self.ptr + offset_bytes, out.nbytes), "D2H")
        return out

    def free(self):
        if self.ptr:
            _capi.synth().synth_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PinnedArray:
    """numpy view over cudaMallocHost memory."""

    def __init__(self, dtype, count):
        self.dtype = np.dtype(dtype)
        self.count = int(count)
        p = C.c_void_p()
        _chk(_capi.synth().synth_malloc_host(C.byref(p), max(self.count * self.dtype.itemsize, 1)), "cudaMallocHost")
        self.ptr = p.value
        buf = (C.c_char * (self.count * self.dtype.itemsize)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=self.count)

    def free(self):
        if self.ptr:
            self.array = None
            _capi.synth().synth_free_host(self.ptr)
            self.ptr = None


class DeviceColumns:
    def __init__(self, n: int):
        self.n = int(n)
        self.bufs = [DeviceBuffer(self.n * 4), DeviceBuffer(self.n * 4), DeviceBuffer(self.n * 8), DeviceBuffer(self.n * 4)]

    def columns(self, offset=0) -> TxColumns:
        b = self.bufs
        return TxColumns(b[0].ptr + 4 * offset, b[1].ptr + 4 * offset, b[2].ptr + 8 * offset, b[3].ptr + 4 * offset)

    def to_host(self):
        b = self.bufs
        return (b[0].download(np.uint32, self.n), b[1].download(np.uint32, self.n), b[2].download(np.int64, self.n),
                b[3].download(np.int32, self.n))

    def free(self):
        for b in self.bufs:
            b.free()


def zipf_cdf(num_services: int, s: float = 1.0, hot=None) -> np.ndarray:
    """P(service <= k) * 2^32 as uint32.  hot=(h, mass): the first h services share `mass` of
    the traffic equally, the rest is Zipf over the remaining services (cfg5's two-tier skew)."""
    w = 1.0 / np.arange(1, num_services + 1, dtype=np.float64) ** s
    if hot:
        h, mass = hot
        w[:h] = mass / h
        if h < num_services:
            w[h:] *= (1.0 - mass) / w[h:].sum()
    c = np.cumsum(w / w.sum())
    cdf = np.minimum(np.floor(c * 4294967296.0), 4294967295.0).astype(np.uint64).astype(np.uint32)
    cdf[-1] = 0xFFFFFFFF
    return cdf


def duration_tables(num_services: int, seed: int = 0, q_size: int = 4096, sigma: float = 0.5):
    rng = np.random.default_rng(seed ^ 0x5EED)
    base = rng.integers(20, 420, size=num_services).astype(np.uint32)
    # lognormal(0, sigma) quantiles * 1024 via the rational approximation of the probit (Acklam)
    p = (np.arange(q_size) + 0.5) / q_size
    z = _probit(p)
    q = np.rint(np.exp(sigma * z) * 1024.0).astype(np.uint32)
    perm = np.random.default_rng(seed ^ 0xBEEF).permutation(q_size)
    return base, q[perm]


def _probit(p):
    a = [-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02, 1.383577518672690e+02,
         -3.066479806614716e+01, 2.506628277459239e+00]
    b = [-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02, 6.680131188771972e+01,
         -1.328068155288572e+01]
    c = [-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00, -2.549732539343734e+00,
         4.374664141464968e+00, 2.938163982698783e+00]
    d = [7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00, 3.754408661907416e+00]
    p = np.asarray(p, dtype=np.float64)
    out = np.empty_like(p)
    lo, hi = p < 0.02425, p > 1 - 0.02425
    mid = ~(lo | hi)
    q = np.sqrt(-2 * np.log(p[lo]))
    out[lo] = (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / \
              ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1)
    q = np.sqrt(-2 * np.log(1 - p[hi]))
    out[hi] = -(((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / \
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1)
    q = p[mid] - 0.5
    r = q * q
    out[mid] = (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q / \
               (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1)
    return out


@dataclass
class StreamSpec:
    seed: int
    n: int
    num_services: int
    num_jvms: int = 1
    t0_ms: int = 1578409200000
    span_ms: int = 3_600_000
    jitter_ms: int = 2000
    elapsed_max: int = 60000
    shift_mod: int = 0
    shift_mul: int = 5
    zipf_s: float = 1.0
    hot: tuple = None


class SynthStream:
    """One synthetic stream; chunks of it can be generated on the device or on the host."""

    def __init__(self, spec: StreamSpec, device: bool = True):
        self.spec = spec
        self.cdf = zipf_cdf(spec.num_services, spec.zipf_s, spec.hot)
        self.base, self.q = duration_tables(spec.num_services, spec.seed)
        self._dev = None
        if device:
            self._dev = [DeviceBuffer(self.cdf.nbytes).upload(self.cdf), DeviceBuffer(self.base.nbytes).upload(self.base),
                         DeviceBuffer(self.q.nbytes).upload(self.q)]

    def _params(self, first, device, t0_ms=None, seed=None) -> SynthParams:
        s = self.spec
        p = SynthParams()
        p.seed, p.n, p.first = (s.seed if seed is None else seed), s.n, first
        p.num_services, p.num_jvms = s.num_services, s.num_jvms
        p.t0_ms, p.span_ms = (s.t0_ms if t0_ms is None else t0_ms), s.span_ms
        p.jitter_ms, p.elapsed_max = s.jitter_ms, s.elapsed_max
        p.shift_mod, p.shift_mul = s.shift_mod, s.shift_mul
        p.q_size = self.q.size
        if device:
            p.cdf, p.base, p.q = (b.ptr for b in self._dev)
        else:
            p.cdf, p.base, p.q = self.cdf.ctypes.data, self.base.ctypes.data, self.q.ctypes.data
        return p

    def generate_device(self, cols: DeviceColumns, first=0, count=None, t0_ms=None, seed=None, offset=0):
        count = self.spec.n - first if count is None else count
        p = self._params(first, True, t0_ms, seed)
        c = cols.columns(offset)
        _chk(_capi.synth().synth_generate_device(C.byref(p), count, C.byref(c)), "synth_generate_device")

    def generate_host(self, first=0, count=None, t0_ms=None, seed=None):
        count = self.spec.n - first if count is None else count
        svc, jvm = np.empty(count, np.uint32), np.empty(count, np.uint32)
        ts, el = np.empty(count, np.int64), np.empty(count, np.int32)
        p = self._params(first, False, t0_ms, seed)
        c = TxColumns(svc.ctypes.data, jvm.ctypes.data, ts.ctypes.data, el.ctypes.data)
        _capi.synth().synth_generate_host(C.byref(p), count, C.byref(c))
        return svc, jvm, ts, el

    def free(self):
        for b in self._dev or []:
            b.free()


def ring_fill_device(buf: DeviceBuffer, rows: int, stream: "SynthStream", seed: int, base_dev=None, num_keys=None):
    """Fills buf as [rows][3][K] with a plausible baseline history for `stream`'s services (or, for a key shard, for
    the `num_keys` local keys whose duration scales are in the device table `base_dev`)."""
    sp = stream.spec
    if base_dev is None:
        K = sp.num_services * sp.num_jvms
        _chk(_capi.synth().synth_ring_fill_device(buf.ptr, rows, 3 * K, seed, stream._dev[1].ptr, sp.num_services, K), "ring_fill")
    else:
        _chk(_capi.synth().synth_ring_fill_device(buf.ptr, rows, 3 * num_keys, seed, base_dev, num_keys, num_keys), "ring_fill")


class ShardPlan:
    """Placement of a global key space on `nranks` ranks: owner[key], the dense id each key has on its owner, and
    per rank the global keys in local-id order.  owner=None: apm_owner_of_key (splitmix64 hash)."""

    def __init__(self, num_keys: int, nranks: int, owner=None):
        from . import sharding
        self.num_keys, self.nranks = int(num_keys), int(nranks)
        self.owner = (sharding.owner_of_keys(np.arange(num_keys, dtype=np.uint64), nranks) if owner is None
                      else np.ascontiguousarray(owner, dtype=np.uint32))
        order = np.argsort(self.owner, kind="stable")
        counts = np.bincount(self.owner, minlength=nranks)
        starts = np.concatenate([[0], np.cumsum(counts)])
        self.local_id = np.empty(num_keys, np.uint32)
        self.local_id[order] = (np.arange(num_keys) - starts[self.owner[order]]).astype(np.uint32)
        self.keys_of = [order[starts[r]:starts[r + 1]] for r in range(nranks)]
        self.counts = counts
        self._dev = None

    def device(self):
        if self._dev is None:
            self._dev = (DeviceBuffer(self.owner.nbytes).upload(self.owner), DeviceBuffer(self.local_id.nbytes).upload(self.local_id))
        return self._dev

    def free(self):
        for b in self._dev or []:
            b.free()
        self._dev = None


def expected_key_counts(stream: "SynthStream") -> np.ndarray:
    """Expected transactions per key (jvm * num_services + service) of the whole stream, from its popularity table."""
    sp = stream.spec
    c = stream.cdf.astype(np.float64)
    p = np.diff(np.concatenate([[-1.0], c])) / 4294967296.0
    return np.tile(p * sp.n / sp.num_jvms, sp.num_jvms)


def generate_shard_device(stream: "SynthStream", plan: ShardPlan, rank: int, cols: DeviceColumns, bucket_ms: int, latest: int,
                          t0_ms=None, seed=None):
    """One rank's shard of the global stream into `cols` (arrival order, dense local ids, jvm 0).  Returns
    (transactions written, tick marks of the GLOBAL stream as (pos in the shard, label), latestBucket afterwards)."""
    S = _capi.synth()
    sp = stream.spec
    n = sp.n
    p = stream._params(0, True, t0_ms, seed)
    tiles = int(S.synth_shard_tiles(n))
    owner_d, local_d = plan.device()
    tc, tm = DeviceBuffer(tiles * 4), DeviceBuffer(tiles * 8)
    _chk(S.synth_shard_count_device(C.byref(p), n, owner_d.ptr, rank, bucket_ms, tc.ptr, tm.ptr), "shard_count")
    cnt = tc.download(np.uint32, tiles).astype(np.uint64)
    mx = tm.download(np.int64, tiles)
    prefix = np.concatenate([[0], np.cumsum(cnt)[:-1]]).astype(np.uint64)
    carry = np.maximum.accumulate(np.concatenate([[latest], mx]))[:-1].astype(np.int64)
    total = int(cnt.sum())
    if total > cols.n:
        raise RuntimeError(f"shard of {total} transactions does not fit the {cols.n} allocated")
    tp, tcar = DeviceBuffer(tiles * 8).upload(prefix), DeviceBuffer(tiles * 8).upload(carry)
    cap = 1 << 14
    marks_d, nm = DeviceBuffer(cap * 16), DeviceBuffer(4).upload(np.zeros(1, np.uint32))
    c = cols.columns()
    _chk(S.synth_shard_emit_device(C.byref(p), n, owner_d.ptr, local_d.ptr, rank, bucket_ms, tp.ptr, tcar.ptr, tm.ptr, C.byref(c),
                                   marks_d.ptr, nm.ptr, cap), "shard_emit")
    k = int(nm.download(np.uint32, 1)[0])
    if k > cap:
        raise RuntimeError("too many ticks in one synthetic step")
    marks = marks_d.download(_capi.MARK_DTYPE, k)
    marks = marks[np.argsort(marks["label"], kind="stable")]
    for b in (tc, tm, tp, tcar, marks_d, nm):
        b.free()
    return total, marks, int(max(latest, mx.max() if tiles else latest))


def ring_fill_host(rows: int, stream: "SynthStream", seed: int, cols) -> np.ndarray:
    """The same values for a selection of columns (series * K + key), as [rows][len(cols)]."""
    sp = stream.spec
    K = sp.num_services * sp.num_jvms
    cols = np.ascontiguousarray(cols, dtype=np.uint64)
    out = np.empty((rows, cols.size), dtype=np.float64)
    _capi.synth().synth_ring_fill_host(out.ctypes.data, rows, 3 * K, seed, stream.base.ctypes.data, sp.num_services, K,
                                       cols.ctypes.data, cols.size)
    return out
