"""CPU-side checks of the C ABI: the library loads, exports every symbol include/apm_b200.h
declares, its pure host helpers agree with the oracle, and compute entry points refuse to run
without a device (no fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from apmbackend_b200 import _capi
from apmbackend_b200.engine import EngineConfig
from oracle import capi as ocapi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "apm_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(apm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(_capi.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in apm_b200.h but not exported"
    assert set(names) == set(_capi.SYMBOLS), set(names) ^ set(_capi.SYMBOLS)


def test_struct_layouts_match_the_header():
    assert C.sizeof(_capi.ApmConfig) == 200 == C.sizeof(ocapi.ApmConfig)
    assert _capi.STAT_DTYPE == ocapi.STAT_DTYPE and _capi.FS_DTYPE == ocapi.FS_DTYPE
    assert C.sizeof(_capi.Counters) == 96 and C.sizeof(_capi.TickResult) == 40


def test_quantize_matches_oracle_to_fixed():
    lib = _capi.lib()
    rng = np.random.default_rng(3)
    xs = np.concatenate([rng.uniform(0, 70000, 20000), rng.integers(0, 1400000, 20000) / 20.0,
                         rng.integers(0, 280000, 5000) / 4.0, rng.integers(0, 560000, 5000) / 8.0,
                         rng.integers(1, 10**7, 20000) / rng.integers(1, 5000, 20000),
                         [0.0, 0.05, 0.15, 0.25, 0.35, 1.45, 1234.05, 8.345, 1e-300, 5e-324, 0.049999999999999996]])
    for d in (1, 2):
        got = np.array([lib.apm_quantize(float(x), d) for x in xs])
        want = np.array([ocapi.quantize(float(x), d) for x in xs])
        assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
    assert np.isnan(lib.apm_quantize(float("nan"), 1))
    assert lib.apm_quantize(-0.04, 1) == ocapi.quantize(-0.04, 1)


def test_owner_of_key_is_balanced_and_stable():
    lib = _capi.lib()
    owners = np.array([lib.apm_owner_of_key(k, 8) for k in range(80000)])
    counts = np.bincount(owners, minlength=8)
    assert counts.min() > 0.95 * 10000 and counts.max() < 1.05 * 10000
    assert lib.apm_owner_of_key(12345, 1) == 0
    assert [lib.apm_owner_of_key(k, 4) for k in (0, 1, 2, 3)] == [lib.apm_owner_of_key(k, 4) for k in (0, 1, 2, 3)]


def test_config_validation_and_no_cpu_fallback():
    lib = _capi.lib()
    h = C.c_void_p()
    bad = EngineConfig(num_services=0).to_c()
    assert lib.apm_ctx_create(C.byref(bad), C.byref(h)) == _capi.E_INVALID
    assert b"num_services" in lib.apm_last_error(None)
    bad = EngineConfig(num_services=4, window=60, buffer=10).to_c()
    assert lib.apm_ctx_create(C.byref(bad), C.byref(h)) == _capi.E_INVALID
    bad = EngineConfig(num_services=4, lags=(360, 100)).to_c()
    assert lib.apm_ctx_create(C.byref(bad), C.byref(h)) == _capi.E_INVALID
    if _capi.device_count() == 0:
        ok = EngineConfig(num_services=4).to_c()
        assert lib.apm_ctx_create(C.byref(ok), C.byref(h)) == _capi.E_DEVICE
        assert b"no CPU fallback" in lib.apm_last_error(None)
        assert not h.value


def test_null_arguments_are_rejected_not_crashed():
    lib = _capi.lib()
    assert lib.apm_tick(None, 5, None) == _capi.E_INVALID
    assert lib.apm_sync(None) == _capi.E_INVALID
    assert lib.apm_fetch_stats(None, None, 0, None) == _capi.E_INVALID
    lib.apm_ctx_destroy(None)
