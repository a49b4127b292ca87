"""Thin object wrapper over the C ABI (include/apm_b200.h): one Engine == one apm_ctx."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

from . import _capi
from ._capi import ApmConfig, ApmError, Counters, TickResult, TxColumns


@dataclass
class EngineConfig:
    """Resolved numeric settings of both stages (config/apm_config.json:105-107, 126-172).
    Defaults are the reference's: 10 s buckets, window 30, buffer 6, p75/p95, lags 360 and 8640."""
    num_services: int
    num_jvms: int = 1
    bucket_ms: int = 10000
    window: int = 30
    buffer: int = 6
    percentiles: Sequence[float] = (75.0, 95.0)
    z_series_pct: Sequence[int] = (0, 1)
    lags: Sequence[int] = (360, 8640)
    thresholds: Sequence[float] = (20.0, 15.0)
    influences: Sequence[float] = (0.1, 0.0)
    flags: int = _capi.F_PERCENTILES | _capi.F_ZSCORE | _capi.F_KEEP_STATS | _capi.F_KEEP_FS | _capi.F_ALERTS
    sd_mode: int = 0
    max_tx_per_bucket: int = 1 << 20
    max_batch: int = 1 << 20
    out_ticks: int = 64
    max_alerts_per_tick: int = 4096
    device: int = 0

    def to_c(self) -> ApmConfig:
        c = ApmConfig()
        c.struct_size = C.sizeof(ApmConfig)
        c.device = self.device
        c.num_services, c.num_jvms = self.num_services, self.num_jvms
        c.bucket_ms, c.window_buckets, c.buffer_buckets = self.bucket_ms, self.window, self.buffer
        c.num_percentiles = len(self.percentiles)
        for i, p in enumerate(self.percentiles):
            c.percentiles[i] = p
        c.z_series_pct[0], c.z_series_pct[1] = self.z_series_pct
        c.num_lags = len(self.lags)
        for i, l in enumerate(self.lags):
            c.lags[i] = l
            c.default_threshold[i] = self.thresholds[i]
            c.default_influence[i] = self.influences[i]
        c.sd_mode, c.flags = self.sd_mode, self.flags
        c.max_tx_per_bucket, c.max_batch = self.max_tx_per_bucket, self.max_batch
        c.out_ticks, c.max_alerts_per_tick = self.out_ticks, self.max_alerts_per_tick
        return c


def _cols(service_id, jvm_id, end_ts_ms, elapsed_ms):
    s = np.ascontiguousarray(service_id, dtype=np.uint32)
    j = np.ascontiguousarray(jvm_id, dtype=np.uint32)
    t = np.ascontiguousarray(end_ts_ms, dtype=np.int64)
    e = np.ascontiguousarray(elapsed_ms, dtype=np.int32)
    if not (s.size == j.size == t.size == e.size):
        raise ValueError("columns differ in length")
    return s, j, t, e


def pack8(service_id, jvm_id, end_ts_ms, elapsed_ms, num_services: int, num_jvms: int, base_ts_ms: int, out=None):
    """Host columns -> the 8-byte form of apm_process_host_packed8: (words u64, escapes).  Rows with an id out of range
    get key 0xFFFF (needs fewer than 65 535 keys), durations that do not fit 24 bits (or are negative) go to the escape
    list.  `out`: optional uint64 array to fill (e.g. page-locked memory)."""
    s, j, t, e = _cols(service_id, jvm_id, end_ts_ms, elapsed_ms)
    K = num_services * num_jvms
    if K > 65536:
        raise ValueError("the 8-byte form carries 16-bit key ids")
    bad = (s >= num_services) | (j >= num_jvms)
    if bad.any() and K > 65535:
        raise ValueError("no spare key code for rows with ids out of range")
    dt = t - np.int64(base_ts_ms)
    if dt.size and (dt.min() < 0 or dt.max() >= (1 << 24)):
        raise ValueError("endTs - base must fit 24 bits (0 .. 4.6 h)")
    key = np.where(bad, 0xFFFF, j.astype(np.uint64) * np.uint64(num_services) + s.astype(np.uint64)).astype(np.uint64)
    esc_rows = np.flatnonzero((e < 0) | (e >= 0xFFFFFF))
    el = e.astype(np.int64)
    el[esc_rows] = 0xFFFFFF
    w = out if out is not None else np.empty(s.size, np.uint64)
    np.left_shift(key, np.uint64(48), out=w)
    w |= dt.astype(np.uint64) << np.uint64(24)
    w |= el.astype(np.uint64)
    esc = np.zeros(esc_rows.size, _capi.ESCAPE_DTYPE)
    esc["index"] = esc_rows
    esc["elapsed_ms"] = e[esc_rows]
    return w, esc


class Engine:
    def __init__(self, cfg: EngineConfig):
        self.cfg = cfg
        self._lib = _capi.lib()
        self._c = cfg.to_c()
        h = C.c_void_p()
        st = self._lib.apm_ctx_create(C.byref(self._c), C.byref(h))
        if st != _capi.OK:
            raise ApmError(st, (self._lib.apm_last_error(None) or b"").decode())
        self._h = h
        self._pinned = {}
        self.K = cfg.num_services * cfg.num_jvms

    # ---- lifecycle
    def close(self):
        if getattr(self, "_h", None):
            for p, _ in getattr(self, "_pinned", {}).values():
                self._lib.apm_host_free(p)
            self._pinned = {}
            self._lib.apm_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, st, res=None):
        if st != _capi.OK:
            err = ApmError(st, (self._lib.apm_last_error(self._h) or b"").decode())
            err.result = res  # APM_E_WOULDBLOCK: what the chunks before the refused one consumed
            raise err

    @property
    def handle(self):
        return self._h

    # ---- ingest / tick
    def process_host(self, service_id, jvm_id, end_ts_ms, elapsed_ms) -> TickResult:
        """A run of consumeMsg calls (stream_calc_stats.js:331-371) over host columns in arrival order."""
        s, j, t, e = _cols(service_id, jvm_id, end_ts_ms, elapsed_ms)
        cols = TxColumns(s.ctypes.data, j.ctypes.data, t.ctypes.data, e.ctypes.data)
        res = TickResult()
        self._check(self._lib.apm_process_host(self._h, C.byref(cols), s.size, C.byref(res)), res)
        return res

    def process_host_packed(self, key, end_ts_delta_ms, elapsed_ms, base_ts_ms: int) -> TickResult:
        """apm_process_host_packed: 12-byte transactions (key = jvm << 16 | service, endTs - base, elapsed), all uint32."""
        k = np.ascontiguousarray(key, dtype=np.uint32)
        t = np.ascontiguousarray(end_ts_delta_ms, dtype=np.uint32)
        e = np.ascontiguousarray(elapsed_ms, dtype=np.uint32)
        if not (k.size == t.size == e.size):
            raise ValueError("columns differ in length")
        cols = _capi.TxPacked(k.ctypes.data, t.ctypes.data, e.ctypes.data)
        res = TickResult()
        self._check(self._lib.apm_process_host_packed(self._h, C.byref(cols), int(base_ts_ms), k.size, C.byref(res)))
        return res

    def process_host_packed8(self, words, escapes, base_ts_ms: int) -> TickResult:
        """apm_process_host_packed8: one u64 per transaction (key:16 | endTs - base:24 | elapsed:24) + the escape list."""
        w = np.ascontiguousarray(words, dtype=np.uint64)
        e = np.ascontiguousarray(escapes, dtype=_capi.ESCAPE_DTYPE) if escapes is not None else np.zeros(0, _capi.ESCAPE_DTYPE)
        res = TickResult()
        self._check(self._lib.apm_process_host_packed8(self._h, w.ctypes.data if w.size else None, e.ctypes.data if e.size else None,
                                                       e.size, int(base_ts_ms), w.size, C.byref(res)))
        return res

    def process_device(self, cols: TxColumns, n: int) -> TickResult:
        res = TickResult()
        self._check(self._lib.apm_process_device(self._h, C.byref(cols), n, C.byref(res)))
        return res

    def set_tick_marks(self, marks):
        """apm_set_tick_marks: the next process_* call ticks where the host says (key-sharded deployments; ticks are
        global, stream_calc_stats.js:348).  marks: structured array (pos u64, label i64) or a sequence of pairs."""
        m = np.ascontiguousarray(marks, dtype=_capi.MARK_DTYPE) if isinstance(marks, np.ndarray) else \
            np.array([tuple(x) for x in marks], dtype=_capi.MARK_DTYPE)
        self._check(self._lib.apm_set_tick_marks(self._h, m.ctypes.data if m.size else None, m.size))

    # ---- log feeder
    @staticmethod
    def name_hash(name) -> int:
        b = name.encode() if isinstance(name, str) else bytes(name)
        return int(_capi.lib().apm_name_hash(b, len(b)))

    def dict_set(self, names_to_ids: dict):
        """Service dictionary of the device parser: rewritten service name -> service id."""
        h = np.array([self.name_hash(k) for k in names_to_ids], np.uint64)
        v = np.array(list(names_to_ids.values()), np.uint32)
        self._check(self._lib.apm_dict_set(self._h, h.ctypes.data, v.ctypes.data, h.size))

    def parse_log(self, text: bytes, file_type: int, jvm_id: int, tz_rules=((-(1 << 62), 0),), want_records=True):
        """apm_parse_log_host: (records in line order, ParseResult); the fed exits stay on the device."""
        tz = np.array(tz_rules, np.int64).reshape(-1, 2)
        buf = text if isinstance(text, np.ndarray) else np.frombuffer(text, np.uint8)  # an ndarray may be pinned memory
        res = _capi.ParseResult()
        n = C.c_uint64(0)
        cap = max(1, int(np.count_nonzero(buf == 10)) + 1) if want_records else 0
        out = np.zeros(cap, _capi.LOG_DTYPE) if want_records else None
        self._check(self._lib.apm_parse_log_host(self._h, buf.ctypes.data if buf.size else None, buf.size, file_type, jvm_id,
                                                 tz.ctypes.data, tz.shape[0], out.ctypes.data if want_records else None, cap,
                                                 C.byref(n), C.byref(res)))
        return (out[:n.value] if want_records else None), res

    def pinned_bytes(self, nbytes: int) -> np.ndarray:
        """A page-locked uint8 array (apm_host_alloc), e.g. for log text that is to be parsed at PCIe speed."""
        return self._out_buffer("text", np.uint8, nbytes, True)[:nbytes]

    def feed_parsed(self) -> TickResult:
        res = TickResult()
        self._check(self._lib.apm_feed_parsed(self._h, C.byref(res)))
        return res

    def add_batch(self, service_id, jvm_id, end_ts_ms, elapsed_ms):
        """addData (:115-134) for a batch that contains no tick (the caller ticks first, as consumeMsg does)."""
        s, j, t, e = _cols(service_id, jvm_id, end_ts_ms, elapsed_ms)
        off = 0
        while off < s.size:
            m = min(s.size - off, self.cfg.max_batch)
            pin = TxColumns()
            self._check(self._lib.apm_ingest_acquire(self._h, m, C.byref(pin)))
            for ptr, arr, dt in ((pin.service_id, s, np.uint32), (pin.jvm_id, j, np.uint32),
                                 (pin.end_ts_ms, t, np.int64), (pin.elapsed_ms, e, np.int32)):
                dst = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(m,))
                dst[:] = arr[off:off + m]
            self._check(self._lib.apm_ingest_commit(self._h, m))
            off += m

    def tick(self, latest_bucket: int) -> TickResult:
        res = TickResult()
        self._check(self._lib.apm_tick(self._h, int(latest_bucket), C.byref(res)))
        return res

    def sync(self):
        self._check(self._lib.apm_sync(self._h))

    # ---- results
    @property
    def ticks_buffered(self) -> int:
        return int(self._lib.apm_ticks_buffered(self._h))

    def _out_buffer(self, kind: str, dtype, cap: int, pinned: bool) -> np.ndarray:
        """Output array for a fetch.  pinned=True reuses one page-locked buffer per record kind
        (apm_host_alloc): the returned view is then only valid until the next fetch of that kind."""
        if not pinned:
            return np.zeros(max(cap, 1), dtype=dtype)
        have = self._pinned.get(kind)
        if have is None or have[1].size < cap:
            if have is not None:
                self._lib.apm_host_free(have[0])
            p = C.c_void_p()
            nbytes = max(cap, 1) * np.dtype(dtype).itemsize
            self._check(self._lib.apm_host_alloc(nbytes, C.byref(p)))
            arr = np.frombuffer((C.c_char * nbytes).from_address(p.value), dtype=dtype, count=max(cap, 1))
            self._pinned[kind] = have = (p, arr)
        return have[1]

    def bind_fs_output(self):
        """fs records stream into one pinned buffer tick by tick; fetch_fs(pinned=True) then returns a view of it."""
        cap = self.cfg.out_ticks * self.K * len(self.cfg.lags)
        out = self._out_buffer("fs", _capi.FS_DTYPE, cap, True)
        self._check(self._lib.apm_bind_fs_output(self._h, out.ctypes.data, out.size))

    def fetch_stats(self, pinned: bool = False) -> np.ndarray:
        cap = self.ticks_buffered * self.K
        out = self._out_buffer("stats", _capi.STAT_DTYPE, cap, pinned)
        n = C.c_uint64(0)
        self._check(self._lib.apm_fetch_stats(self._h, out.ctypes.data, cap, C.byref(n)))
        return out[:n.value]

    def fetch_fs(self, pinned: bool = False) -> np.ndarray:
        cap = self.ticks_buffered * self.K * len(self.cfg.lags)
        out = self._out_buffer("fs", _capi.FS_DTYPE, cap, pinned)
        n = C.c_uint64(0)
        self._check(self._lib.apm_fetch_fs(self._h, out.ctypes.data, max(cap, 0), C.byref(n)))
        return out[:n.value]

    def fetch_alerts(self, pinned: bool = False) -> np.ndarray:
        cap = self.ticks_buffered * self.cfg.max_alerts_per_tick
        out = self._out_buffer("alerts", _capi.FS_DTYPE, cap, pinned)
        n = C.c_uint64(0)
        self._check(self._lib.apm_fetch_alerts(self._h, out.ctypes.data, cap, C.byref(n)))
        return out[:n.value]

    def fetch_passthrough(self, cap: int = 1 << 22):
        """(end_ts, seq, per-tick counts) of the transactions the retained ticks released, oldest first."""
        ts = np.empty(cap, np.int64)
        seq = np.empty(cap, np.uint32)
        counts = np.zeros(max(self.cfg.out_ticks, 1), np.uint32)
        n, nt = C.c_uint64(0), C.c_uint32(0)
        self._check(self._lib.apm_fetch_passthrough(self._h, ts.ctypes.data, seq.ctypes.data, cap, C.byref(n), counts.ctypes.data,
                                                    counts.size, C.byref(nt)))
        return ts[:n.value], seq[:n.value], counts[:nt.value]

    def release_ticks(self):
        self._check(self._lib.apm_release_ticks(self._h))

    # ---- z-score settings and lag lists
    def set_key_params(self, key_lo: int, lag_idx: int, thr, infl):
        t = np.ascontiguousarray(thr, dtype=np.float64)
        f = np.ascontiguousarray(infl, dtype=np.float64)
        self._check(self._lib.apm_set_key_params(self._h, key_lo, t.size, lag_idx, t.ctypes.data, f.ctypes.data))

    def set_alert_rules(self, window=60, required=45, alert_on_both_only=True, suppressed_lags=(), hard_max_ms=10000.0,
                        hard_min_ms=200.0, hard_min_tpm=1.0, enabled=True):
        """The counter core of processFSEntry (stream_process_alerts.js:348-434) behind every tick;
        defaults are the reference's (config/apm_config.json:189-204)."""
        r = _capi.AlertRules()
        r.struct_size = C.sizeof(_capi.AlertRules)
        r.enabled, r.window, r.required, r.alert_on_both_only = int(enabled), window, required, int(alert_on_both_only)
        r.suppressed_lag_mask = sum(1 << list(self.cfg.lags).index(l) for l in suppressed_lags if l in self.cfg.lags)
        r.hard_max_ms, r.hard_min_ms, r.hard_min_tpm = hard_max_ms, hard_min_ms, hard_min_tpm
        self._check(self._lib.apm_set_alert_rules(self._h, C.byref(r)))

    def set_alert_key_params(self, key_lo: int, hard_max_ms, suppressed):
        h = np.ascontiguousarray(hard_max_ms, dtype=np.float64)
        s = np.ascontiguousarray(suppressed, dtype=np.uint8)
        assert h.size == s.size
        self._check(self._lib.apm_set_alert_key_params(self._h, key_lo, h.size, h.ctypes.data, s.ctypes.data))

    def drop_lag(self, lag_idx: int):
        self._check(self._lib.apm_drop_lag(self._h, lag_idx))

    def ring_load(self, key: int, lag_idx: int, series: int, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        self._check(self._lib.apm_ring_load(self._h, key, lag_idx, series, v.ctypes.data, v.size))

    def ring_load_all(self, lag_idx: int, values=None, length: int = 0, device_ptr: Optional[int] = None):
        if device_ptr is not None:
            self._check(self._lib.apm_ring_load_all(self._h, lag_idx, C.c_void_p(device_ptr), length, 1))
        else:
            v = np.ascontiguousarray(values, dtype=np.float64)
            assert v.shape[1:] == (3, self.K)
            self._check(self._lib.apm_ring_load_all(self._h, lag_idx, v.ctypes.data, v.shape[0], 0))

    def ring_read(self, key: int, lag_idx: int, series: int) -> np.ndarray:
        cap = int(self.cfg.lags[lag_idx])
        out = np.empty(cap, dtype=np.float64)
        n = C.c_uint32(0)
        self._check(self._lib.apm_ring_read(self._h, key, lag_idx, series, out.ctypes.data, cap, C.byref(n)))
        return out[:n.value].copy()

    # ---- interop with the reference's JSON resume files
    @property
    def latest(self) -> int:
        return int(self._lib.apm_latest(self._h))

    def set_latest(self, latest_bucket: int):
        self._check(self._lib.apm_set_latest(self._h, int(latest_bucket)))

    def mark_seen(self, key_lo: int, n: int = 1):
        self._check(self._lib.apm_mark_seen(self._h, key_lo, n))

    def log_read(self, label: int) -> np.ndarray:
        """(key << 32 | elapsed) words of one retained bucket (arrival order until the bucket is sorted)."""
        n = C.c_uint64(0)
        self._check(self._lib.apm_log_read(self._h, int(label), None, 0, C.byref(n)))
        out = np.empty(max(n.value, 1), dtype=np.uint64)
        if n.value:
            self._check(self._lib.apm_log_read(self._h, int(label), out.ctypes.data, out.size, C.byref(n)))
        return out[:n.value]

    # ---- checkpoint
    def state_save(self, path: str):
        self._check(self._lib.apm_state_save(self._h, path.encode()))

    def state_load(self, path: str):
        self._check(self._lib.apm_state_load(self._h, path.encode()))

    def counters(self) -> Counters:
        c = Counters()
        self._check(self._lib.apm_get_counters(self._h, C.byref(c)))
        return c

    def set_overlap(self, on: bool):
        self._check(self._lib.apm_set_overlap(self._h, 1 if on else 0))

    def timer_start(self):
        self._check(self._lib.apm_timer_start(self._h))

    def timer_stop(self) -> float:
        ms = C.c_double(0)
        self._check(self._lib.apm_timer_stop(self._h, C.byref(ms)))
        return ms.value

    def profile_enable(self, on: bool = True):
        self._check(self._lib.apm_profile_enable(self._h, 1 if on else 0))

    def profile_read(self) -> dict:
        kt = _capi.KernelTimes()
        self._check(self._lib.apm_profile_read(self._h, C.byref(kt)))
        return {g: {"ms": kt.ms[i], "spans": int(kt.spans[i])} for i, g in enumerate(_capi.KERNEL_GROUPS)}

    def reset_counters(self):
        self._check(self._lib.apm_reset_counters(self._h))

    # ---- multi-GPU
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        st = _capi.lib().apm_comm_unique_id(buf)
        if st != _capi.OK:
            raise ApmError(st, (_capi.lib().apm_last_error(None) or b"").decode())
        return buf.raw

    def comm_init(self, rank: int, nranks: int, unique_id: bytes):
        self._check(self._lib.apm_comm_init(self._h, rank, nranks, unique_id))

    def comm_set_capacity(self, records_per_block: int):
        self._check(self._lib.apm_comm_set_capacity(self._h, int(records_per_block)))

    def bind_gather_output(self, cap: int):
        """Every tick's alerts are all-gathered as the tick finishes; gather_alerts(cap, pinned=True) then returns them."""
        out = self._out_buffer("gather", _capi.FS_DTYPE, cap, True)
        self._check(self._lib.apm_bind_gather_output(self._h, out.ctypes.data, out.size))

    def gather_alerts(self, cap: int, pinned: bool = False) -> np.ndarray:
        out = self._out_buffer("gather", _capi.FS_DTYPE, cap, pinned)
        n = C.c_uint64(0)
        self._check(self._lib.apm_gather_alerts(self._h, out.ctypes.data, cap, C.byref(n)))
        return out[:n.value]
