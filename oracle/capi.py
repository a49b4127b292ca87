"""
ctypes binding of oracle/liboracle.so (the C++ restatement in apm_oracle.cpp).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this.  The product package
never does; apmbackend_b200 fails loudly without its CUDA library instead.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

STAT_DTYPE = np.dtype([
    ("tick_ts", "<i8"), ("key", "<u4"), ("flags", "<u4"), ("count", "<u8"), ("sum", "<u8"),
    ("tpm", "<f8"), ("avg", "<f8"), ("pct", "<f8", (4,)), ("min", "<i4"), ("max", "<i4"), ("stddev", "<f8"),
])
FS_DTYPE = np.dtype([
    ("tick_ts", "<i8"), ("key", "<u4"), ("lag", "<u4"), ("tpm", "<f8"),
    ("val", "<f8", (3,)), ("avg", "<f8", (3,)), ("lb", "<f8", (3,)), ("ub", "<f8", (3,)),
    ("signal", "i1", (3,)), ("flags", "u1"), ("pad", "u1", (4,)),
])
assert STAT_DTYPE.itemsize == 96 and FS_DTYPE.itemsize == 128


class ApmConfig(C.Structure):
    """Mirror of apm_config (include/apm_b200.h)."""
    _fields_ = [
        ("struct_size", C.c_uint32), ("device", C.c_int32),
        ("num_services", C.c_uint32), ("num_jvms", C.c_uint32),
        ("bucket_ms", C.c_int64),
        ("window_buckets", C.c_int32), ("buffer_buckets", C.c_int32),
        ("num_percentiles", C.c_int32), ("z_series_pct", C.c_int32 * 2),
        ("percentiles", C.c_double * 4),
        ("num_lags", C.c_int32), ("lags", C.c_int32 * 4),
        ("default_threshold", C.c_double * 4), ("default_influence", C.c_double * 4),
        ("sd_mode", C.c_int32), ("flags", C.c_uint32),
        ("max_tx_per_bucket", C.c_uint64), ("max_batch", C.c_uint64),
        ("out_ticks", C.c_uint32), ("max_alerts_per_tick", C.c_uint32),
    ]


class AlertRules(C.Structure):
    """Mirror of apm_alert_rules (include/apm_b200.h); defaults = config/apm_config.json:189-204."""
    _fields_ = [("struct_size", C.c_uint32), ("enabled", C.c_int32), ("window", C.c_int32), ("required", C.c_int32),
                ("alert_on_both_only", C.c_int32), ("suppressed_lag_mask", C.c_uint32),
                ("hard_max_ms", C.c_double), ("hard_min_ms", C.c_double), ("hard_min_tpm", C.c_double)]


def make_alert_rules(window=60, required=45, alert_on_both_only=True, suppressed_lag_mask=0, hard_max_ms=10000.0,
                     hard_min_ms=200.0, hard_min_tpm=1.0, enabled=True) -> AlertRules:
    r = AlertRules()
    r.struct_size = C.sizeof(AlertRules)
    r.enabled, r.window, r.required = int(enabled), window, required
    r.alert_on_both_only, r.suppressed_lag_mask = int(alert_on_both_only), suppressed_lag_mask
    r.hard_max_ms, r.hard_min_ms, r.hard_min_tpm = hard_max_ms, hard_min_ms, hard_min_tpm
    return r


F_PERCENTILES, F_ZSCORE, F_KEEP_STATS, F_KEEP_FS, F_ALERTS = 1, 2, 4, 8, 16


def make_config(num_services, num_jvms=1, bucket_ms=10000, window=30, buffer=6, percentiles=(75.0, 95.0),
                z_series_pct=(0, 1), lags=(360, 8640), thresholds=(20.0, 15.0), influences=(0.1, 0.0),
                flags=F_PERCENTILES | F_ZSCORE | F_KEEP_STATS | F_KEEP_FS | F_ALERTS, sd_mode=0,
                max_tx_per_bucket=1 << 20, max_batch=1 << 20, out_ticks=64, max_alerts_per_tick=4096,
                device=0) -> ApmConfig:
    c = ApmConfig()
    c.struct_size = C.sizeof(ApmConfig)
    c.device = device
    c.num_services, c.num_jvms = num_services, num_jvms
    c.bucket_ms, c.window_buckets, c.buffer_buckets = bucket_ms, window, buffer
    c.num_percentiles = len(percentiles)
    for i, p in enumerate(percentiles):
        c.percentiles[i] = p
    c.z_series_pct[0], c.z_series_pct[1] = z_series_pct
    c.num_lags = len(lags)
    for i, l in enumerate(lags):
        c.lags[i] = l
        c.default_threshold[i] = thresholds[i]
        c.default_influence[i] = influences[i]
    c.sd_mode, c.flags = sd_mode, flags
    c.max_tx_per_bucket, c.max_batch = max_tx_per_bucket, max_batch
    c.out_ticks, c.max_alerts_per_tick = out_ticks, max_alerts_per_tick
    return c


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "apm_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_quantize.restype = C.c_double
        L.orc_quantize.argtypes = [C.c_double, C.c_int]
        L.orc_average.restype = C.c_double
        L.orc_average.argtypes = [C.c_void_p, C.c_uint64]
        L.orc_standard_deviation.restype = C.c_double
        L.orc_standard_deviation.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_int)]
        L.orc_calc_percentile.restype = C.c_double
        L.orc_calc_percentile.argtypes = [C.c_void_p, C.c_uint64, C.c_double]
        L.orc_process_zscore.restype = C.c_int
        L.orc_process_zscore.argtypes = [C.c_int64, C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_uint64,
                                         C.POINTER(C.c_double * 4)]
        L.orc_to_fixed.argtypes = [C.c_double, C.c_int, C.c_char_p, C.c_int]
        L.orc_new.restype = C.c_void_p
        L.orc_new.argtypes = [C.POINTER(ApmConfig), C.c_int]
        for name in ("orc_free", "orc_clear_outputs"):
            getattr(L, name).argtypes = [C.c_void_p]
            getattr(L, name).restype = None
        L.orc_set_filter.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_keep_outputs.argtypes = [C.c_void_p, C.c_int]
        L.orc_set_key_params.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_double, C.c_double]
        L.orc_ring_load.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_uint32]
        L.orc_ring_read.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_uint32]
        L.orc_ring_read.restype = C.c_uint32
        L.orc_set_alert_rules.argtypes = [C.c_void_p, C.POINTER(AlertRules)]
        L.orc_set_alert_key_params.argtypes = [C.c_void_p, C.c_uint32, C.c_double, C.c_int]
        L.orc_process_fs_entry.argtypes = [C.POINTER(AlertRules), C.c_int, C.c_int, C.c_double, C.c_void_p,
                                           C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.orc_consume.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        for name in ("orc_num_stats", "orc_num_fs", "orc_ticks", "orc_ingested"):
            getattr(L, name).argtypes = [C.c_void_p]
            getattr(L, name).restype = C.c_uint64
        L.orc_latest.argtypes = [C.c_void_p]
        L.orc_latest.restype = C.c_int64
        L.orc_copy_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_copy_fs.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_format_st.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.orc_format_fs.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
        _LIB = L
    return _LIB


def to_fixed(x: float, digits: int) -> str:
    buf = C.create_string_buffer(128)
    lib().orc_to_fixed(x, digits, buf, 128)
    return buf.value.decode()


def quantize(x: float, digits: int) -> float:
    return lib().orc_quantize(x, digits)


def calc_percentile(sorted_vals, p: float) -> float:
    a = np.ascontiguousarray(sorted_vals, dtype=np.int32)
    return lib().orc_calc_percentile(a.ctypes.data, a.size, float(p))


def average(lst) -> float:
    a = np.ascontiguousarray(lst, dtype=np.float64)
    return lib().orc_average(a.ctypes.data, a.size)


def standard_deviation(lst):
    a = np.ascontiguousarray(lst, dtype=np.float64)
    u = C.c_int(0)
    r = lib().orc_standard_deviation(a.ctypes.data, a.size, C.byref(u))
    return None if u.value else r


def process_zscore(lag, thr, infl, new_value, lst):
    a = np.ascontiguousarray(lst, dtype=np.float64)
    out = (C.c_double * 4)()
    sig = lib().orc_process_zscore(int(lag), thr, infl, new_value, a.ctypes.data, a.size, C.byref(out))
    return [out[0], out[1], out[2], out[3], sig]


class Pipeline:
    """Both stages back to back, as the `stats` queue connects them."""

    def __init__(self, cfg: ApmConfig, fast: bool = True):
        self.cfg = cfg
        self._h = lib().orc_new(C.byref(cfg), 1 if fast else 0)

    def close(self):
        if self._h:
            lib().orc_free(self._h)
            self._h = None

    __del__ = close

    def set_filter(self, keep: np.ndarray):
        keep = np.ascontiguousarray(keep, dtype=np.uint8)
        assert keep.size == self.cfg.num_services * self.cfg.num_jvms
        lib().orc_set_filter(self._h, keep.ctypes.data)

    def keep_outputs(self, keep: bool):
        lib().orc_keep_outputs(self._h, 1 if keep else 0)

    def set_key_params(self, key, lag_idx, thr, infl):
        lib().orc_set_key_params(self._h, key, lag_idx, thr, infl)

    def ring_load(self, key, lag_idx, series, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        lib().orc_ring_load(self._h, key, lag_idx, series, v.ctypes.data, v.size)

    def ring_read(self, key, lag_idx, series):
        cap = int(self.cfg.lags[lag_idx])
        out = np.empty(cap, dtype=np.float64)
        n = lib().orc_ring_read(self._h, key, lag_idx, series, out.ctypes.data, cap)
        return out[:n].copy()

    def set_alert_rules(self, rules: AlertRules):
        lib().orc_set_alert_rules(self._h, C.byref(rules))

    def set_alert_key_params(self, key, hard_max_ms=0.0, suppressed=False):
        lib().orc_set_alert_key_params(self._h, key, hard_max_ms, int(suppressed))

    def consume(self, service_id, jvm_id, end_ts_ms, elapsed_ms):
        s = np.ascontiguousarray(service_id, dtype=np.uint32)
        j = np.ascontiguousarray(jvm_id, dtype=np.uint32)
        t = np.ascontiguousarray(end_ts_ms, dtype=np.int64)
        e = np.ascontiguousarray(elapsed_ms, dtype=np.int32)
        assert s.size == j.size == t.size == e.size
        lib().orc_consume(self._h, s.ctypes.data, j.ctypes.data, t.ctypes.data, e.ctypes.data, s.size)

    def stats(self) -> np.ndarray:
        n = lib().orc_num_stats(self._h)
        out = np.zeros(n, dtype=STAT_DTYPE)
        if n:
            lib().orc_copy_stats(self._h, out.ctypes.data)
        return out

    def fs(self) -> np.ndarray:
        n = lib().orc_num_fs(self._h)
        out = np.zeros(n, dtype=FS_DTYPE)
        if n:
            lib().orc_copy_fs(self._h, out.ctypes.data)
        return out

    def clear_outputs(self):
        lib().orc_clear_outputs(self._h)

    @property
    def ticks(self):
        return lib().orc_ticks(self._h)

    @property
    def latest(self):
        return lib().orc_latest(self._h)

    @property
    def ingested(self):
        return lib().orc_ingested(self._h)


def format_st(rec, server: str, service: str, pa=0, pb=1) -> str:
    r = np.ascontiguousarray(rec).reshape(1)
    buf = C.create_string_buffer(512)
    lib().orc_format_st(r.ctypes.data, server.encode(), service.encode(), pa, pb, buf, 512)
    return buf.value.decode()


def format_fs(rec, server: str, service: str) -> str:
    r = np.ascontiguousarray(rec).reshape(1)
    buf = C.create_string_buffer(1024)
    lib().orc_format_fs(r.ctypes.data, server.encode(), service.encode(), buf, 1024)
    return buf.value.decode()


def process_fs_entry(rules: AlertRules, rec, count: int, lag_suppressed=False, service_suppressed=False, hard_max_override=0.0):
    """One call of processFSEntry's counter core: returns (triggerAlert, causes bits, new count)."""
    r = np.ascontiguousarray(rec).reshape(1)
    c, causes = C.c_uint32(count), C.c_uint32(0)
    t = lib().orc_process_fs_entry(C.byref(rules), int(lag_suppressed), int(service_suppressed), hard_max_override,
                                   r.ctypes.data, C.byref(c), C.byref(causes))
    return bool(t), causes.value, c.value
