"""ctypes binding of libapm_b200.so (include/apm_b200.h).  Fails loudly when the
CUDA library is missing - there is deliberately no other implementation to fall
back to."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libapm_b200.so")
SYNTH_PATH = os.path.join(_HERE, "libapm_synth.so")

OK, E_INVALID, E_DEVICE, E_CAPACITY, E_WOULDBLOCK, E_STATE, E_IO, E_COMM = range(8)
F_PERCENTILES, F_ZSCORE, F_KEEP_STATS, F_KEEP_FS, F_ALERTS, F_INCREMENTAL, F_PASSTHROUGH = 1, 2, 4, 8, 16, 32, 64
MAX_PCT, MAX_LAGS, NUM_SERIES = 4, 4, 3

STAT_DTYPE = np.dtype([
    ("tick_ts", "<i8"), ("key", "<u4"), ("flags", "<u4"), ("count", "<u8"), ("sum", "<u8"),
    ("tpm", "<f8"), ("avg", "<f8"), ("pct", "<f8", (4,)), ("min", "<i4"), ("max", "<i4"), ("stddev", "<f8"),
])
FS_DTYPE = np.dtype([
    ("tick_ts", "<i8"), ("key", "<u4"), ("lag", "<u4"), ("tpm", "<f8"),
    ("val", "<f8", (3,)), ("avg", "<f8", (3,)), ("lb", "<f8", (3,)), ("ub", "<f8", (3,)),
    ("signal", "i1", (3,)), ("flags", "u1"), ("pad", "u1", (4,)),
])
LOG_DTYPE = np.dtype([
    ("line_offset", "<u8"), ("line_length", "<u4"), ("kind", "u1"), ("flags", "u1"), ("service_length", "<u2"),
    ("service_offset", "<u4"), ("service_id", "<u4"), ("elapsed_ms", "<i4"), ("logid_length", "<u4"),
    ("ts_ms", "<i8"), ("start_ms", "<i8"), ("logid_hash", "<u8"), ("name_hash", "<u8"),
])
assert STAT_DTYPE.itemsize == 96 and FS_DTYPE.itemsize == 128 and LOG_DTYPE.itemsize == 64
LK_CT_START, LK_CT_STOP, LK_EJB_ENTRY, LK_EJB_EXIT = 1, 2, 3, 4
LR_LOGID_EMPTY, LR_BAD_TIME, LR_BAD_ELAPSED, LR_PAIRED, LR_OPEN_START, LR_BAF_INFO, LR_TOP_LEVEL, LR_FED = 1, 2, 4, 8, 16, 32, 64, 128
NO_SERVICE = 0xFFFFFFFF


class ApmConfig(C.Structure):
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


class TxColumns(C.Structure):
    _fields_ = [("service_id", C.c_void_p), ("jvm_id", C.c_void_p), ("end_ts_ms", C.c_void_p),
                ("elapsed_ms", C.c_void_p)]


class TickResult(C.Structure):
    _fields_ = [("ticks", C.c_uint64), ("tx_ingested", C.c_uint64), ("tx_dropped", C.c_uint64),
                ("tx_invalid", C.c_uint64), ("latest_bucket", C.c_int64)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "tx_ingested", "tx_dropped", "tx_invalid", "ticks", "stat_records", "fs_records", "alerts",
        "h2d_bytes", "d2h_bytes", "kernel_launches", "keys_seen", "inc_fallbacks")]


class AlertRules(C.Structure):
    """apm_alert_rules; defaults are config/apm_config.json:189-204."""
    _fields_ = [("struct_size", C.c_uint32), ("enabled", C.c_int32), ("window", C.c_int32), ("required", C.c_int32),
                ("alert_on_both_only", C.c_int32), ("suppressed_lag_mask", C.c_uint32),
                ("hard_max_ms", C.c_double), ("hard_min_ms", C.c_double), ("hard_min_tpm", C.c_double)]


CAUSE_AVG_HARD_MAX, CAUSE_P75_HARD_MAX, CAUSE_AVG_UB, CAUSE_P75_UB, CAUSE_BOTH_UB = 1, 2, 4, 8, 16


class KernelTimes(C.Structure):
    _fields_ = [("ms", C.c_double * 12), ("spans", C.c_uint64 * 12)]


KERNEL_GROUPS = ("tick_finder", "keyed_reduce", "slot_clear", "radix_sort", "window_percentiles", "rolling_baseline",
                 "alert_compaction", "list_realign", "passthrough")


class ApmError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"apm_b200 status {status}: {msg}")
        self.status = status


_lib = None
_synth = None

# every symbol include/apm_b200.h declares (tests check the library exports them all)
class TxPacked(C.Structure):
    _fields_ = [("key", C.c_void_p), ("end_ts_delta_ms", C.c_void_p), ("elapsed_ms", C.c_void_p)]


class TickMark(C.Structure):
    """apm_tick_mark: the tick for bucket `label` fires before transaction `pos` of the next process_* call."""
    _fields_ = [("pos", C.c_uint64), ("label", C.c_int64)]


MARK_DTYPE = np.dtype([("pos", "<u8"), ("label", "<i8")])
ESCAPE_DTYPE = np.dtype([("index", "<u4"), ("elapsed_ms", "<i4")])


class PartitionOut(C.Structure):
    _fields_ = [("cols", TxColumns), ("cap", C.c_uint64), ("n", C.c_uint64), ("marks", C.c_void_p),
                ("marks_cap", C.c_uint32), ("num_marks", C.c_uint32)]


class ParseResult(C.Structure):
    _fields_ = [("lines", C.c_uint64), ("records", C.c_uint64), ("exits", C.c_uint64), ("fed", C.c_uint64),
                ("unknown_service", C.c_uint64), ("malformed", C.c_uint64)]


SYMBOLS = {
    "apm_abi_version": (C.c_uint32, []),
    "apm_ctx_create": (C.c_int32, [C.POINTER(ApmConfig), C.POINTER(C.c_void_p)]),
    "apm_ctx_destroy": (None, [C.c_void_p]),
    "apm_last_error": (C.c_char_p, [C.c_void_p]),
    "apm_ingest_acquire": (C.c_int32, [C.c_void_p, C.c_uint64, C.POINTER(TxColumns)]),
    "apm_ingest_commit": (C.c_int32, [C.c_void_p, C.c_uint64]),
    "apm_tick": (C.c_int32, [C.c_void_p, C.c_int64, C.POINTER(TickResult)]),
    "apm_process_host": (C.c_int32, [C.c_void_p, C.POINTER(TxColumns), C.c_uint64, C.POINTER(TickResult)]),
    "apm_process_device": (C.c_int32, [C.c_void_p, C.POINTER(TxColumns), C.c_uint64, C.POINTER(TickResult)]),
    "apm_set_key_params": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int32, C.c_void_p, C.c_void_p]),
    "apm_drop_lag": (C.c_int32, [C.c_void_p, C.c_int32]),
    "apm_set_alert_rules": (C.c_int32, [C.c_void_p, C.POINTER(AlertRules)]),
    "apm_set_alert_key_params": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]),
    "apm_fetch_stats": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "apm_fetch_fs": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "apm_fetch_alerts": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "apm_release_ticks": (C.c_int32, [C.c_void_p]),
    "apm_fetch_passthrough": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.c_void_p,
                                          C.c_uint32, C.POINTER(C.c_uint32)]),
    "apm_ticks_buffered": (C.c_uint32, [C.c_void_p]),
    "apm_host_alloc": (C.c_int32, [C.c_uint64, C.POINTER(C.c_void_p)]),
    "apm_host_free": (None, [C.c_void_p]),
    "apm_bind_fs_output": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "apm_sync": (C.c_int32, [C.c_void_p]),
    "apm_ring_load": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_int32, C.c_int32, C.c_void_p, C.c_uint32]),
    "apm_ring_read": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_int32, C.c_int32, C.c_void_p, C.c_uint32,
                                  C.POINTER(C.c_uint32)]),
    "apm_ring_load_all": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_uint32, C.c_int32]),
    "apm_state_save": (C.c_int32, [C.c_void_p, C.c_char_p]),
    "apm_state_load": (C.c_int32, [C.c_void_p, C.c_char_p]),
    "apm_profile_enable": (C.c_int32, [C.c_void_p, C.c_int32]),
    "apm_profile_read": (C.c_int32, [C.c_void_p, C.POINTER(KernelTimes)]),
    "apm_set_overlap": (C.c_int32, [C.c_void_p, C.c_int32]),
    "apm_timer_start": (C.c_int32, [C.c_void_p]),
    "apm_timer_stop": (C.c_int32, [C.c_void_p, C.POINTER(C.c_double)]),
    "apm_set_latest": (C.c_int32, [C.c_void_p, C.c_int64]),
    "apm_mark_seen": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_uint32]),
    "apm_log_read": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "apm_latest": (C.c_int64, [C.c_void_p]),
    "apm_get_counters": (C.c_int32, [C.c_void_p, C.POINTER(Counters)]),
    "apm_reset_counters": (C.c_int32, [C.c_void_p]),
    "apm_quantize": (C.c_double, [C.c_double, C.c_int32]),
    "apm_sorted_percentiles": (C.c_int32, [C.c_int32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int32, C.c_void_p,
                                           C.c_void_p]),
    "apm_sorted_percentiles_host": (C.c_int32, [C.c_int32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int32, C.c_void_p,
                                                C.c_void_p]),
    "apm_sorted_percentiles_noalloc": (C.c_int32, [C.c_int32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int32,
                                                   C.c_void_p, C.c_void_p]),
    "apm_owner_of_key": (C.c_uint32, [C.c_uint64, C.c_uint32]),
    "apm_owner_table_build": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint64, C.c_void_p]),
    "apm_owner_table_build_capped": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_double, C.c_void_p]),
    "apm_comm_unique_id": (C.c_int32, [C.c_void_p]),
    "apm_comm_init": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "apm_gather_alerts": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "apm_process_host_packed": (C.c_int32, [C.c_void_p, C.POINTER(TxPacked), C.c_int64, C.c_uint64, C.POINTER(TickResult)]),
    "apm_name_hash": (C.c_uint64, [C.c_char_p, C.c_uint64]),
    "apm_dict_set": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "apm_parse_log_host": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32,
                                       C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(ParseResult)]),
    "apm_feed_parsed": (C.c_int32, [C.c_void_p, C.POINTER(TickResult)]),
    "apm_process_host_packed8": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int64, C.c_uint64,
                                             C.POINTER(TickResult)]),
    "apm_preload_kernels": (C.c_int32, [C.c_int32]),
    "apm_comm_set_capacity": (C.c_int32, [C.c_void_p, C.c_uint32]),
    "apm_bind_gather_output": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "apm_set_tick_marks": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint32]),
    "apm_partition_host": (C.c_int32, [C.POINTER(TxColumns), C.c_uint64, C.c_int64, C.c_int64, C.c_uint32, C.c_uint32,
                                       C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]),
}


def lib():
    """The CUDA library.  Raises if it has not been built - never substitutes anything."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C apmbackend_b200/csrc`.  apmbackend_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError here = library out of date
            fn.restype = res
            fn.argtypes = args
        if L.apm_abi_version() != 1:
            raise ImportError("libapm_b200.so ABI version mismatch")
        _lib = L
    return _lib


def synth():
    """Bench / test support library (synthetic streams, raw device buffers)."""
    global _synth
    if _synth is None:
        if not os.path.exists(SYNTH_PATH):
            raise ImportError(f"{SYNTH_PATH} is missing: run `make -C apmbackend_b200/csrc`")
        S = C.CDLL(SYNTH_PATH)
        S.synth_device_count.restype = C.c_int
        S.synth_malloc.argtypes = [C.POINTER(C.c_void_p), C.c_uint64]
        S.synth_malloc_host.argtypes = [C.POINTER(C.c_void_p), C.c_uint64]
        S.synth_free.argtypes = [C.c_void_p]
        S.synth_free_host.argtypes = [C.c_void_p]
        S.synth_h2d.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        S.synth_d2h.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        S.synth_memset.argtypes = [C.c_void_p, C.c_int, C.c_uint64]
        S.synth_error_string.restype = C.c_char_p
        S.synth_error_string.argtypes = [C.c_int]
        S.synth_generate_device.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(TxColumns)]
        S.synth_generate_host.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(TxColumns)]
        S.synth_generate_host.restype = None
        S.synth_ring_fill_device.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint32, C.c_uint32]
        S.synth_ring_fill_host.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint32, C.c_uint32,
                                           C.c_void_p, C.c_uint64]
        S.synth_ring_fill_host.restype = None
        S.synth_mem_info.argtypes = [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        S.synth_shard_tiles.argtypes = [C.c_uint64]
        S.synth_shard_tiles.restype = C.c_int
        S.synth_shard_count_device.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_int64, C.c_void_p, C.c_void_p]
        S.synth_shard_emit_device.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int64, C.c_void_p,
                                              C.c_void_p, C.c_void_p, C.POINTER(TxColumns), C.c_void_p, C.c_void_p, C.c_uint32]
        _synth = S
    return _synth


def device_count() -> int:
    try:
        return int(synth().synth_device_count())
    except Exception:
        return 0
