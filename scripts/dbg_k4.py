"""Debug build only (make -C apmbackend_b200/csrc clean all NVCCFLAGS+=-DAPM_K4_DEBUG): where the adders of
baseline_few_kernel spend their cycles on the reference's own shape (400 keys, lags 360 + 8640, lists full)."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, ".")
from apmbackend_b200 import Engine, EngineConfig, _capi  # noqa: E402
from tests.streams import make_stream  # noqa: E402

S, J, lags = 200, 2, (360, 8640)
K = S * J
eng = Engine(EngineConfig(num_services=S, num_jvms=J, bucket_ms=10000, window=30, buffer=6, percentiles=(75.0, 95.0), lags=lags,
                          thresholds=(20.0, 15.0), influences=(0.1, 0.0), max_batch=1 << 16, max_tx_per_bucket=1 << 14, out_ticks=64))
rng = np.random.default_rng(5)
for li, L in enumerate(lags):
    eng.ring_load_all(li, np.round(rng.uniform(50, 400, (L, 3, K)), 1))
svc, jvm, ts, el = make_stream(9, 30000, S, J, 10000, 60)
eng.process_host(svc, jvm, ts, el)
eng.fetch_fs()
eng.release_ticks()
out = (C.c_ulonglong * 32)()
lib = _capi.lib()
rc = lib.apm_debug_k4(out)
names = ["prologue", "wait stage", "lds", "wait other adder", "chain", "post+release", "-", "-", "total", "chunks"]
for w in range(2):
    print("adder", w, {n: int(out[w * 16 + i]) for i, n in enumerate(names) if n != "-"}, "rc", rc)
eng.close()
